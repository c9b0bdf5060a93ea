// Mask ingest: the step before the hot path.  The reference loads every
// silhouette with Image1b::Load (image.h:76-101: stb_image decode of one file, single-threaded, then
// a copy into the image) before it can call Carve; at 360 views of 2048 x 2048 that serial decode
// takes longer than the whole GPU pass.  Here a batch of files is decoded by a pool of host threads
// straight into pinned memory and handed to the carver, whose copy stream overlaps the uploads with
// the kernels.
//
// Formats: 8-bit greyscale PNG, non-interlaced (what the reference's data/mask_*.png are) and
// binary PGM (P5, maxval 255).  Anything else is refused with VCB_ERR_UNSUPPORTED - there is no
// silent conversion; such files stay with the reference's own loader.
//
// The PNG path is self-contained: chunk walk, zlib/DEFLATE decoder (RFC 1950 / 1951: stored, fixed
// and dynamic Huffman blocks, canonical codes decoded by length counts) and the five scanline
// filters of the PNG specification (None, Sub, Up, Average, Paeth).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <thread>

#include "carver_internal.cuh"

using namespace vcb;

namespace {

// ---- DEFLATE -------------------------------------------------------------------------------------
struct BitReader {
  const uint8_t* p;
  size_t n, pos = 0;
  uint32_t buf = 0;
  int cnt = 0;
  bool bad = false;
  int bits(int need) {
    while (cnt < need) {
      if (pos >= n) {
        bad = true;
        return 0;
      }
      buf |= (uint32_t)p[pos++] << cnt;
      cnt += 8;
    }
    const int v = (int)(buf & ((1u << need) - 1u));
    buf >>= need;
    cnt -= need;
    return v;
  }
};

struct Huffman {  // canonical code: number of codes per length + symbols ordered by code
  uint16_t count[16];
  uint16_t symbol[288];
  bool build(const uint8_t* lengths, int n) {
    memset(count, 0, sizeof(count));
    for (int i = 0; i < n; i++) count[lengths[i]]++;
    count[0] = 0;
    int left = 1;
    for (int len = 1; len < 16; len++) {
      left = (left << 1) - count[len];
      if (left < 0) return false;  // over-subscribed
    }
    uint16_t offs[16];
    offs[1] = 0;
    for (int len = 1; len < 15; len++) offs[len + 1] = offs[len] + count[len];
    for (int i = 0; i < n; i++) {
      if (lengths[i]) symbol[offs[lengths[i]]++] = (uint16_t)i;
    }
    return true;
  }
  int decode(BitReader& br) const {
    int code = 0, first = 0, index = 0;
    for (int len = 1; len < 16; len++) {
      code |= br.bits(1);
      if (br.bad) return -1;
      const int c = count[len];
      if (code - c < first) return symbol[index + (code - first)];
      index += c;
      first = (first + c) << 1;
      code <<= 1;
    }
    return -1;
  }
};

const uint16_t kLenBase[29] = {3,  4,  5,  6,  7,  8,  9,  10, 11,  13,  15,  17,  19,  23, 27,
                               31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const uint8_t kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2,
                               2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const uint16_t kDistBase[30] = {1,   2,   3,   4,   5,   7,    9,    13,   17,   25,
                                33,  49,  65,  97,  129, 193,  257,  385,  513,  769,
                                1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const uint8_t kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2,  3,  3,  4,  4,  5,  5,  6,
                                6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

// inflates a raw DEFLATE stream into out[0..cap); returns the number of bytes or -1
int64_t inflate_raw(const uint8_t* src, size_t n, uint8_t* out, size_t cap) {
  BitReader br{src, n};
  size_t o = 0;
  Huffman lit, dist;
  for (;;) {
    const int last = br.bits(1), type = br.bits(2);
    if (br.bad) return -1;
    if (type == 0) {  // stored
      br.buf = 0;
      br.cnt = 0;
      if (br.pos + 4 > n) return -1;
      const unsigned len = src[br.pos] | (src[br.pos + 1] << 8);
      const unsigned nlen = src[br.pos + 2] | (src[br.pos + 3] << 8);
      br.pos += 4;
      if ((len ^ 0xffffu) != nlen || br.pos + len > n || o + len > cap) return -1;
      memcpy(out + o, src + br.pos, len);
      br.pos += len;
      o += len;
    } else if (type == 1 || type == 2) {
      uint8_t lengths[320];
      if (type == 1) {
        int i = 0;
        for (; i < 144; i++) lengths[i] = 8;
        for (; i < 256; i++) lengths[i] = 9;
        for (; i < 280; i++) lengths[i] = 7;
        for (; i < 288; i++) lengths[i] = 8;
        lit.build(lengths, 288);
        for (i = 0; i < 30; i++) lengths[i] = 5;
        dist.build(lengths, 30);
      } else {
        const int nlen = br.bits(5) + 257, ndist = br.bits(5) + 1, ncode = br.bits(4) + 4;
        if (br.bad || nlen > 286 || ndist > 30) return -1;
        static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
        uint8_t cl[19] = {0};
        for (int i = 0; i < ncode; i++) cl[order[i]] = (uint8_t)br.bits(3);
        Huffman lencode;
        if (!lencode.build(cl, 19)) return -1;
        int i = 0;
        while (i < nlen + ndist) {
          const int sym = lencode.decode(br);
          if (sym < 0) return -1;
          if (sym < 16) {
            lengths[i++] = (uint8_t)sym;
          } else {
            int rep, val = 0;
            if (sym == 16) {
              if (i == 0) return -1;
              val = lengths[i - 1];
              rep = 3 + br.bits(2);
            } else if (sym == 17) {
              rep = 3 + br.bits(3);
            } else {
              rep = 11 + br.bits(7);
            }
            if (br.bad || i + rep > nlen + ndist) return -1;
            while (rep--) lengths[i++] = (uint8_t)val;
          }
        }
        if (lengths[256] == 0) return -1;
        if (!lit.build(lengths, nlen) || !dist.build(lengths + nlen, ndist)) return -1;
      }
      for (;;) {
        const int sym = lit.decode(br);
        if (sym < 0) return -1;
        if (sym < 256) {
          if (o >= cap) return -1;
          out[o++] = (uint8_t)sym;
        } else if (sym == 256) {
          break;
        } else {
          const int li = sym - 257;
          if (li >= 29) return -1;
          const size_t len = kLenBase[li] + (size_t)br.bits(kLenExtra[li]);
          const int di = dist.decode(br);
          if (di < 0 || di >= 30) return -1;
          const size_t d = kDistBase[di] + (size_t)br.bits(kDistExtra[di]);
          if (br.bad || d > o || o + len > cap) return -1;
          if (d >= len) {
            memcpy(out + o, out + o - d, len);
          } else {
            for (size_t k = 0; k < len; k++) out[o + k] = out[o + k - d];  // overlapping run
          }
          o += len;
        }
      }
    } else {
      return -1;
    }
    if (last) break;
  }
  return (int64_t)o;
}

uint32_t be32(const uint8_t* p) {
  return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}

int paeth(int a, int b, int c) {
  const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

}  // namespace

namespace {

struct FileBytes {
  std::vector<uint8_t> data;
  bool read(const char* path) {
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    const long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    data.resize(n > 0 ? (size_t)n : 0);
    const bool ok = n >= 0 && fread(data.data(), 1, data.size(), f) == data.size();
    fclose(f);
    return ok;
  }
};

// Decodes one file.  With pixels == nullptr only the size is reported.  Returns VCB_OK or an error
// code with a message in `err`.
int decode_mask(const char* path, uint8_t* pixels, size_t capacity, int32_t* width, int32_t* height,
                std::string* err) {
  FileBytes fb;
  if (!fb.read(path)) {
    *err = std::string("cannot read ") + path;
    return VCB_ERR_INVALID;
  }
  const std::vector<uint8_t>& d = fb.data;
  static const uint8_t kPngSig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  if (d.size() >= 8 && memcmp(d.data(), kPngSig, 8) == 0) {
    int w = 0, h = 0;
    std::vector<uint8_t> idat;
    bool header = false, end = false;
    for (size_t p = 8; p + 12 <= d.size() && !end;) {
      const uint32_t len = be32(&d[p]);
      const uint8_t* type = &d[p + 4];
      if (p + 12 + (size_t)len > d.size()) break;
      const uint8_t* body = &d[p + 8];
      if (memcmp(type, "IHDR", 4) == 0 && len == 13) {
        w = (int)be32(body);
        h = (int)be32(body + 4);
        if (body[8] != 8 || body[9] != 0 || body[12] != 0 || body[10] != 0 || body[11] != 0) {
          *err = std::string(path) + ": only 8-bit greyscale, non-interlaced PNG is decoded here "
                 "(use the reference's Image1b::Load for other kinds)";
          return VCB_ERR_UNSUPPORTED;
        }
        header = true;
      } else if (memcmp(type, "IDAT", 4) == 0) {
        idat.insert(idat.end(), body, body + len);
      } else if (memcmp(type, "IEND", 4) == 0) {
        end = true;
      }
      p += 12 + (size_t)len;
    }
    if (!header || w < 1 || h < 1 || w > kMaxImageDim || h > kMaxImageDim || idat.size() < 6) {
      *err = std::string(path) + ": damaged PNG";
      return VCB_ERR_INVALID;
    }
    *width = w;
    *height = h;
    if (!pixels) return VCB_OK;
    if (capacity < (size_t)w * h) {
      *err = std::string(path) + ": image larger than the buffer";
      return VCB_ERR_INVALID;
    }
    // zlib wrapper (RFC 1950): 2 header bytes, DEFLATE data, 4 bytes of Adler-32 (not verified)
    if ((idat[0] & 0x0f) != 8 || ((idat[0] << 8) | idat[1]) % 31 != 0 || (idat[1] & 0x20)) {
      *err = std::string(path) + ": bad zlib header";
      return VCB_ERR_INVALID;
    }
    const size_t stride = (size_t)w + 1;  // filter byte + row
    std::vector<uint8_t> raw(stride * h);
    if (inflate_raw(idat.data() + 2, idat.size() - 2, raw.data(), raw.size()) != (int64_t)raw.size()) {
      *err = std::string(path) + ": damaged PNG data";
      return VCB_ERR_INVALID;
    }
    for (int y = 0; y < h; y++) {  // scanline filters, PNG specification section 9 (bpp = 1)
      const uint8_t* in = &raw[stride * y + 1];
      uint8_t* out = pixels + (size_t)w * y;
      const uint8_t* up = y > 0 ? out - w : nullptr;
      switch (raw[stride * y]) {
        case 0:
          memcpy(out, in, w);
          break;
        case 1:
          for (int x = 0; x < w; x++) out[x] = (uint8_t)(in[x] + (x > 0 ? out[x - 1] : 0));
          break;
        case 2:
          for (int x = 0; x < w; x++) out[x] = (uint8_t)(in[x] + (up ? up[x] : 0));
          break;
        case 3:
          for (int x = 0; x < w; x++)
            out[x] = (uint8_t)(in[x] + (((x > 0 ? out[x - 1] : 0) + (up ? up[x] : 0)) >> 1));
          break;
        case 4:
          for (int x = 0; x < w; x++)
            out[x] = (uint8_t)(in[x] + paeth(x > 0 ? out[x - 1] : 0, up ? up[x] : 0,
                                             (x > 0 && up) ? up[x - 1] : 0));
          break;
        default:
          *err = std::string(path) + ": unknown PNG filter";
          return VCB_ERR_INVALID;
      }
    }
    return VCB_OK;
  }
  if (d.size() >= 2 && d[0] == 'P' && d[1] == '5') {  // binary PGM: P5 <w> <h> <maxval> <one space> data
    size_t p = 2;
    long v[3];
    for (int k = 0; k < 3; k++) {
      for (;;) {  // whitespace and comments
        while (p < d.size() && (d[p] == ' ' || d[p] == '\t' || d[p] == '\n' || d[p] == '\r')) p++;
        if (p < d.size() && d[p] == '#') {
          while (p < d.size() && d[p] != '\n') p++;
        } else {
          break;
        }
      }
      long x = 0;
      size_t digits = 0;
      while (p < d.size() && d[p] >= '0' && d[p] <= '9' && digits < 9) {
        x = x * 10 + (d[p++] - '0');
        digits++;
      }
      if (digits == 0) {
        *err = std::string(path) + ": damaged PGM header";
        return VCB_ERR_INVALID;
      }
      v[k] = x;
    }
    p++;  // the single whitespace byte after maxval
    if (v[2] != 255) {
      *err = std::string(path) + ": only PGM files with maxval 255 are decoded here";
      return VCB_ERR_UNSUPPORTED;
    }
    if (v[0] < 1 || v[1] < 1 || v[0] > kMaxImageDim || v[1] > kMaxImageDim ||
        p + (size_t)v[0] * v[1] > d.size()) {
      *err = std::string(path) + ": damaged PGM";
      return VCB_ERR_INVALID;
    }
    *width = (int32_t)v[0];
    *height = (int32_t)v[1];
    if (!pixels) return VCB_OK;
    if (capacity < (size_t)v[0] * v[1]) {
      *err = std::string(path) + ": image larger than the buffer";
      return VCB_ERR_INVALID;
    }
    memcpy(pixels, &d[p], (size_t)v[0] * v[1]);
    return VCB_OK;
  }
  *err = std::string(path) + ": neither an 8-bit grey PNG nor a binary PGM";
  return VCB_ERR_UNSUPPORTED;
}

}  // namespace

extern "C" {

int vcb_load_mask_file(const char* path, uint8_t* pixels, size_t capacity, int32_t* width,
                       int32_t* height) {
  if (!path || !width || !height) return fail(nullptr, VCB_ERR_INVALID, "null argument");
  std::string err;
  const int rc = decode_mask(path, pixels, capacity, width, height, &err);
  return rc == VCB_OK ? VCB_OK : fail(nullptr, rc, "%s", err.c_str());
}

}  // extern "C"

namespace vcb {

cudaError_t PinnedBuffer::reserve(size_t n) {
  if (n <= bytes) return cudaSuccess;
  release();
  cudaError_t e = cudaMallocHost(&ptr, n);
  if (e == cudaSuccess) bytes = n;
  return e;
}

void PinnedBuffer::release() {
  if (ptr) cudaFreeHost(ptr);
  ptr = nullptr;
  bytes = 0;
}

// Decodes files [0, n) whose `wanted` flag is set (all when wanted == nullptr) into
// out + i * stride, on up to n_threads host threads (<= 0: one per hardware thread); every image
// must have the size its view announces.
int decode_mask_batch(const char* const* paths, const vcb_view* views, int n, const bool* wanted,
                      uint8_t* out, size_t stride, int n_threads, std::string* err) {
  unsigned hw = std::thread::hardware_concurrency();
  if (hw == 0) hw = 4;
  int threads = n_threads > 0 ? n_threads : (int)hw;
  if (threads > n) threads = n;
  std::atomic<int> next(0), status(VCB_OK);
  std::vector<std::string> errors(threads > 0 ? threads : 1);
  auto work = [&](int t) {
    for (;;) {
      const int i = next.fetch_add(1);
      if (i >= n || status.load() != VCB_OK) return;
      if (wanted && !wanted[i]) continue;
      int32_t w = 0, h = 0;
      std::string e;
      int rc = decode_mask(paths[i], out + (size_t)i * stride, stride, &w, &h, &e);
      if (rc == VCB_OK && (w != views[i].width || h != views[i].height)) {
        rc = VCB_ERR_INVALID;
        e = std::string(paths[i]) + ": image is " + std::to_string(w) + " x " + std::to_string(h) +
            ", the view says " + std::to_string(views[i].width) + " x " +
            std::to_string(views[i].height);
      }
      if (rc != VCB_OK) {
        errors[t] = e;
        int expected = VCB_OK;
        status.compare_exchange_strong(expected, rc);
        return;
      }
    }
  };
  if (threads <= 1) {
    work(0);
  } else {
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++) pool.emplace_back(work, t);
    for (auto& th : pool) th.join();
  }
  if (status.load() != VCB_OK) {
    for (auto& e : errors)
      if (!e.empty()) *err = e;
  }
  return status.load();
}

}  // namespace vcb

extern "C" int vcb_carve_mask_files(vcb_carver* c, int32_t n_views, const vcb_view* views,
                                    const char* const* paths, int32_t n_threads) {
  if (!c) return fail(nullptr, VCB_ERR_INVALID, "null carver");
  if (n_views < 0 || (n_views > 0 && (!views || !paths))) return fail(c, VCB_ERR_INVALID, "null argument");
  size_t stride = 0;
  for (int i = 0; i < n_views; i++) {
    int rc = check_view(c, views[i]);
    if (rc != VCB_OK) return rc;
    if (!paths[i]) return fail(c, VCB_ERR_INVALID, "path %d is null", i);
    stride = std::max(stride, ((size_t)views[i].width * views[i].height + 255) & ~(size_t)255);
  }
  int rc = use_device(c, c->opt.device);
  if (rc != VCB_OK) return rc;
  // one launch worth of files at a time: while the GPU works on a batch (the carve call returns as
  // soon as its uploads are done) the host threads decode the next one
  const int batch = kMaxViewsPerLaunch;
  CU(c, c->ingest_pin.reserve(stride * (size_t)std::min(batch, std::max(n_views, 1))));
  for (int first = 0; first < n_views; first += batch) {
    const int nb = std::min(batch, n_views - first);
    std::string err;
    rc = decode_mask_batch(paths + first, views + first, nb, nullptr,
                           static_cast<uint8_t*>(c->ingest_pin.ptr), stride, n_threads, &err);
    if (rc != VCB_OK) return fail(c, rc, "%s", err.c_str());
    std::vector<const uint8_t*> masks(nb);
    for (int i = 0; i < nb; i++) masks[i] = static_cast<uint8_t*>(c->ingest_pin.ptr) + (size_t)i * stride;
    if ((rc = vcb_carve_masks(c, nb, views + first, masks.data(), nullptr)) != VCB_OK) return rc;
  }
  return VCB_OK;
}
