// Dev / test tool: checks vacancy_b200/csrc/ply_format.cuh (the formatting the PLY kernels run on
// the GPU) against snprintf("%g") / snprintf("%d") on the CPU, over every binary exponent with random
// and adversarial mantissas.  Build: g++ -O2 -fopenmp -o ply_format_check tools/ply_format_check.cc
// Prints the number of values checked and of mismatches; exit code 1 on any mismatch.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../vacancy_b200/csrc/ply_format.cuh"

static uint64_t rng_next(uint64_t* s) {
  *s ^= *s << 13;
  *s ^= *s >> 7;
  *s ^= *s << 17;
  return *s;
}

static int check_float(uint32_t bits, long long* bad) {
  float f;
  memcpy(&f, &bits, 4);
  char want[64], got[64];
  const int nw = snprintf(want, sizeof(want), "%g", (double)f);
  const int ng = vcb::ply_format_g6(bits, got);
  if (nw != ng || memcmp(want, got, (size_t)nw) != 0) {
    if (*bad < 10) {
      got[ng < 63 ? ng : 63] = 0;
      fprintf(stderr, "mismatch bits=%08x want=%s got=%s\n", bits, want, got);
    }
    return 1;
  }
  return 0;
}

int main(int argc, char** argv) {
  // "all": every one of the 2^32 bit patterns (a few minutes on 8 cores)
  const bool all = argc > 1 && strcmp(argv[1], "all") == 0;
  const long long per_exp = all ? 0 : (argc > 1 ? atoll(argv[1]) : 100000);
  long long checked = 0, bad = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : checked, bad)
  for (int expo = 0; expo < 256; expo++) {
    uint64_t s = 0x9E3779B97F4A7C15ull * (uint64_t)(expo + 1);
    long long lb = 0;
    for (int sign = 0; sign < 2; sign++) {
      const uint32_t base = ((uint32_t)sign << 31) | ((uint32_t)expo << 23);
      for (long long i = 0; i < per_exp; i++) {  // random mantissas
        lb += check_float(base | (uint32_t)(rng_next(&s) & 0x7fffffu), &bad);
        checked++;
      }
      if (all) {
        for (uint32_t mnt = 0; mnt < (1u << 23); mnt++) lb += check_float(base | mnt, &bad);
        checked += 1ll << 23;
        continue;
      }
      for (uint32_t mnt = 0; mnt < 4096; mnt++) {  // sparse mantissas: few set bits, both ends
        lb += check_float(base | mnt, &bad);
        lb += check_float(base | (mnt << 11), &bad);
        lb += check_float(base | (0x7fffffu - mnt), &bad);
        checked += 3;
      }
    }
    bad += lb;
  }
  // decimal neighbourhoods: floats next to k * 10^j and to the 6-digit rounding boundaries
  for (int j = -44; j <= 38; j++) {
    for (int k = 1; k <= 2000; k++) {
      const double centre[3] = {k * 0.5, 99999.5 + k, 999990.0 + k * 0.005};
      for (int c = 0; c < 3; c++) {
        char buf[64];
        snprintf(buf, sizeof(buf), "%.9ge%d", centre[c], j);
        float f = strtof(buf, nullptr);
        uint32_t bits;
        memcpy(&bits, &f, 4);
        for (int d = -3; d <= 3; d++) {
          bad += check_float(bits + (uint32_t)d, &bad);
          checked++;
        }
      }
    }
  }
  // integers
  uint64_t s = 12345;
  for (long long i = 0; i < 2000000; i++) {
    int32_t v = (int32_t)(uint32_t)rng_next(&s);
    if (i < 64) v = (int32_t)((i & 1) ? -(1ll << (i / 2)) : (1ll << (i / 2)) - 1);
    char want[32], got[32];
    const int nw = snprintf(want, sizeof(want), "%d", v);
    const int ng = vcb::ply_format_int(v, got);
    if (nw != ng || memcmp(want, got, (size_t)nw) != 0) bad++;
    checked++;
  }
  printf("checked %lld mismatches %lld\n", checked, bad);
  return bad != 0;
}
