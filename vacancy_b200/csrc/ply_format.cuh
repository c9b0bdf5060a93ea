// Text formatting of the ASCII PLY writer (F2): byte-compatible with what the reference's
// Mesh::WritePly (mesh.cc:583-631) produces through `std::ofstream << float` / `<< int`, i.e.
// printf("%g") with the default precision of 6 significant digits and printf("%d").
//
// The decimal digits are computed with exact integer arithmetic (no floating point at all), so the
// rounding is the correctly rounded, ties-to-even result glibc prints:
//   |f| = m * 2^e (m < 2^24).  With X the decimal exponent, the six digits are
//   D = round_half_even(m * 2^e * 10^(5 - X)):
//     5 - X >= 0 : m is multiplied by ten (5 - X) times in a 256-bit integer, then shifted right by
//                  -e with the dropped bits deciding the rounding;
//     5 - X <  0 : m * 2^e (< 2^128) is divided by ten (X - 5) times, remainders deciding it.
//   X is taken from the binary exponent (floor(e2 * log10(2)) is X or X - 1) and bumped when the
//   digits overflow.
// The functions are plain C++ (`VCB_HD` = __host__ __device__ under nvcc) so that the same code can be
// checked against snprintf on the CPU by tools/ply_format_check.cc (all 2^32 float bit patterns:
// 0 mismatches); the library only calls them from kernels.
#pragma once

#include <stdint.h>

#ifdef __CUDACC__
#define VCB_HD __host__ __device__ __forceinline__
#else
#define VCB_HD inline
#endif

namespace vcb {

constexpr int kPlyMaxFloatChars = 16;  // "-1.23457e-38" is 12
constexpr int kPlyMaxIntChars = 12;    // "-2147483648" is 11

struct U256 {
  uint32_t w[8];  // little endian limbs
};

VCB_HD void u256_mul10(U256& a) {
  uint64_t carry = 0;
  for (int i = 0; i < 8; i++) {
    const uint64_t t = (uint64_t)a.w[i] * 10u + carry;
    a.w[i] = (uint32_t)t;
    carry = t >> 32;
  }
}
// a /= 10, returns the remainder
VCB_HD uint32_t u256_div10(U256& a) {
  uint64_t rem = 0;
  for (int i = 7; i >= 0; i--) {
    const uint64_t cur = (rem << 32) | a.w[i];
    a.w[i] = (uint32_t)(cur / 10u);
    rem = cur % 10u;
  }
  return (uint32_t)rem;
}
VCB_HD bool u256_bit(const U256& a, int b) { return b < 256 && ((a.w[b >> 5] >> (b & 31)) & 1u); }
// any bit below position b set
VCB_HD bool u256_any_below(const U256& a, int b) {
  for (int i = 0; i < 8; i++) {
    const int lo = i * 32;
    if (b <= lo) break;
    const uint32_t mask = (b - lo >= 32) ? 0xffffffffu : ((1u << (b - lo)) - 1u);
    if (a.w[i] & mask) return true;
  }
  return false;
}
// low 64 bits of (a >> s), limb-wise (s >= 0)
VCB_HD uint64_t u256_shr64(const U256& a, int s) {
  const int limb = s >> 5, off = s & 31;
  uint32_t x[4];
  for (int i = 0; i < 4; i++) x[i] = (limb + i < 8) ? a.w[limb + i] : 0u;
  const uint64_t lo = (uint64_t)x[0] | ((uint64_t)x[1] << 32);
  const uint64_t hi = (uint64_t)x[2] | ((uint64_t)x[3] << 32);
  return off == 0 ? lo : ((lo >> off) | (hi << (64 - off)));
}

// round_half_even(m * 2^e * 10^p) for the six-digit window; the result fits 64 bits when the
// caller's exponent estimate is right or one too small (then it is >= 10^6 and the caller retries)
VCB_HD uint64_t ply_scaled_digits(uint32_t m, int e, int p) {
  U256 a;
  for (int i = 0; i < 8; i++) a.w[i] = 0;
  if (p >= 0) {
    a.w[0] = m;
    for (int i = 0; i < p; i++) u256_mul10(a);
    if (e >= 0) {  // cannot happen for six-digit windows (|f| >= 2^23 has p < 0); kept for safety
      uint64_t v = (uint64_t)a.w[0] | ((uint64_t)a.w[1] << 32);
      return e < 40 ? (v << e) : ~(uint64_t)0;
    }
    const int s = -e;
    uint64_t q = u256_shr64(a, s);
    const bool half = u256_bit(a, s - 1);
    const bool sticky = u256_any_below(a, s - 1);
    if (half && (sticky || (q & 1u))) q++;
    return q;
  }
  // p < 0: (m << e) / 10^(-p); m * 2^e < 2^128 for every finite float; negative e only for
  // 10^6 <= |f| < 2^24, where the shifted-out bits join the remainder test
  bool sticky = false;
  if (e >= 0) {
    const int limb = e >> 5, off = e & 31;
    const uint64_t v = (uint64_t)m << off;
    a.w[limb] = (uint32_t)v;
    if (limb + 1 < 8) a.w[limb + 1] = (uint32_t)(v >> 32);
  } else {
    a.w[0] = m >> (-e);
    sticky = (m & ((1u << (-e)) - 1u)) != 0;
  }
  uint32_t last = 0;
  for (int i = 0; i < -p; i++) {
    sticky = sticky || last != 0;
    last = u256_div10(a);
  }
  uint64_t q = (uint64_t)a.w[0] | ((uint64_t)a.w[1] << 32);
  if (last > 5 || (last == 5 && (sticky || (q & 1u)))) q++;
  return q;
}

VCB_HD int ply_put_uint(uint32_t v, char* out) {
  char tmp[10];
  int n = 0;
  do {
    tmp[n++] = (char)('0' + v % 10u);
    v /= 10u;
  } while (v);
  for (int i = 0; i < n; i++) out[i] = tmp[n - 1 - i];
  return n;
}

// printf("%d")
VCB_HD int ply_format_int(int32_t v, char* out) {
  int n = 0;
  uint32_t u = (uint32_t)v;
  if (v < 0) {
    out[n++] = '-';
    u = 0u - u;
  }
  return n + ply_put_uint(u, out + n);
}

// printf("%g") of a float (promoted to double, precision 6); returns the number of characters
VCB_HD int ply_format_g6(uint32_t bits, char* out) {
  int n = 0;
  if (bits >> 31) out[n++] = '-';
  const uint32_t expo = (bits >> 23) & 0xffu;
  uint32_t frac = bits & 0x7fffffu;
  if (expo == 0xffu) {
    const char* s = frac ? "nan" : "inf";
    for (int i = 0; i < 3; i++) out[n++] = s[i];
    return n;
  }
  if (expo == 0 && frac == 0) {
    out[n++] = '0';
    return n;
  }
  uint32_t m;
  int e;
  if (expo == 0) {  // denormal
    m = frac;
    e = -149;
  } else {
    m = frac | 0x800000u;
    e = (int)expo - 150;
  }
  int top = 31;
  while (!((m >> top) & 1u)) top--;
  const int e2 = top + e;  // floor(log2 |f|)
  int X = (e2 * 78913) >> 18;  // floor(e2 * log10(2)): the decimal exponent or one below it
  uint64_t D = ply_scaled_digits(m, e, 5 - X);
  while (D >= 1000000u) {  // at most once: the estimate is never more than one below
    X++;
    D = ply_scaled_digits(m, e, 5 - X);
  }
  // the six digits, most significant first, and how many remain after stripping trailing zeros
  char d[6];
  uint32_t t = (uint32_t)D;
  for (int i = 5; i >= 0; i--) {
    d[i] = (char)('0' + t % 10u);
    t /= 10u;
  }
  int nd = 6;
  while (nd > 1 && d[nd - 1] == '0') nd--;
  if (X < -4 || X >= 6) {  // exponential style
    out[n++] = d[0];
    if (nd > 1) {
      out[n++] = '.';
      for (int i = 1; i < nd; i++) out[n++] = d[i];
    }
    out[n++] = 'e';
    out[n++] = X < 0 ? '-' : '+';
    const uint32_t ax = (uint32_t)(X < 0 ? -X : X);
    if (ax < 10u) out[n++] = '0';
    n += ply_put_uint(ax, out + n);
  } else if (X >= 0) {  // X + 1 integer digits
    for (int i = 0; i <= X; i++) out[n++] = i < nd ? d[i] : '0';
    if (nd > X + 1) {
      out[n++] = '.';
      for (int i = X + 1; i < nd; i++) out[n++] = d[i];
    }
  } else {  // 0.000ddd
    out[n++] = '0';
    out[n++] = '.';
    for (int i = 0; i < -X - 1; i++) out[n++] = '0';
    for (int i = 0; i < nd; i++) out[n++] = d[i];
  }
  return n;
}

// one vertex line of Mesh::WritePly: "x y z \n" (mesh.cc:616-624, no vertex colours)
constexpr int kPlyMaxVertexLine = 3 * (kPlyMaxFloatChars + 1) + 1;
VCB_HD int ply_vertex_line(uint32_t x, uint32_t y, uint32_t z, char* out) {
  int n = ply_format_g6(x, out);
  out[n++] = ' ';
  n += ply_format_g6(y, out + n);
  out[n++] = ' ';
  n += ply_format_g6(z, out + n);
  out[n++] = ' ';
  out[n++] = '\n';
  return n;
}
// one face line: "3 i j k \n" (mesh.cc:626-629)
constexpr int kPlyMaxFaceLine = 2 + 3 * (kPlyMaxIntChars + 1) + 1;
VCB_HD int ply_face_line(int32_t a, int32_t b, int32_t c, char* out) {
  int n = 0;
  out[n++] = '3';
  out[n++] = ' ';
  n += ply_format_int(a, out + n);
  out[n++] = ' ';
  n += ply_format_int(b, out + n);
  out[n++] = ' ';
  n += ply_format_int(c, out + n);
  out[n++] = ' ';
  out[n++] = '\n';
  return n;
}

}  // namespace vcb
