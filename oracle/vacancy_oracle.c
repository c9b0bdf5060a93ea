/* TEST INFRASTRUCTURE ONLY - see vacancy_oracle.h. Plain-C restatement of the reference hot path.
 * Build: gcc -std=c11 -O2 -ffp-contract=off (no FMA contraction: the reference is built without
 * -march, so x86-64 baseline SSE2 arithmetic, one rounding per operation). */
#include "vacancy_oracle.h"

#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "mc_tables_packed.h"

#ifdef _OPENMP
#include <omp.h>
#endif

/* threads of the one parallel loop (the fusion loop, like voxel_carver.cc:439-442); n < 0 = all
 * cores, 0 = only report. torch.distributed.run exports OMP_NUM_THREADS=1, so the timed baseline
 * sets the count itself. */
int vo_set_threads(int n) {
#ifdef _OPENMP
  if (n < 0) n = omp_get_num_procs();
  if (n > 0) omp_set_num_threads(n);
  return omp_get_max_threads();
#else
  (void)n;
  return 1;
#endif
}

static const uint64_t kTriPacked[256] = VCB_MC_TRI_PACKED_INIT;

float vo_invalid_sdf(void) { return -FLT_MAX; } /* numeric_limits<float>::lowest() */

/* ---- camera ------------------------------------------------------------------------------- */

/* Eigen 3.4.0 LU/InverseImpl.h:125-137 cofactor_3x3<i,j> on a row-major 4x4 (linear part) */
static double cof3(const double* m, int i, int j) {
  int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
  return m[i1 * 4 + j1] * m[i2 * 4 + j2] - m[i1 * 4 + j2] * m[i2 * 4 + j1];
}

/* camera.cc:39-42: w2c_ = c2w_.inverse() (Affine3d => Transform.h:1236-1247: linear().inverse()
 * via InverseImpl.h:161-176, translation = -inv * t), then voxel_carver.cc:438 cast<float>(). */
void vo_w2c_from_c2w(const double* m, float* w2c12) {
  double c0[3], inv[9], det, invdet;
  int r, c;
  c0[0] = cof3(m, 0, 0);
  c0[1] = cof3(m, 1, 0);
  c0[2] = cof3(m, 2, 0);
  /* (cofactors_col0.cwiseProduct(matrix.col(0))).sum(): unrolled redux of a 3-vector */
  det = (c0[0] * m[0] + c0[1] * m[4]) + c0[2] * m[8];
  invdet = 1.0 / det;
  inv[0] = c0[0] * invdet;
  inv[1] = c0[1] * invdet;
  inv[2] = c0[2] * invdet;
  inv[3] = cof3(m, 0, 1) * invdet;
  inv[4] = cof3(m, 1, 1) * invdet;
  inv[5] = cof3(m, 2, 1) * invdet;
  inv[6] = cof3(m, 0, 2) * invdet;
  inv[7] = cof3(m, 1, 2) * invdet;
  inv[8] = cof3(m, 2, 2) * invdet;
  for (r = 0; r < 3; r++) {
    /* -L * t: the negation is applied to the matrix coefficients, then a lazy 3x3 * 3x1 product */
    double t = ((-inv[r * 3 + 0]) * m[3] + (-inv[r * 3 + 1]) * m[7]) + (-inv[r * 3 + 2]) * m[11];
    for (c = 0; c < 3; c++) w2c12[r * 4 + c] = (float)inv[r * 3 + c];
    w2c12[r * 4 + 3] = (float)t;
  }
}

/* ---- 2D signed distance field --------------------------------------------------------------- */

#define AT(img, w, x, y) ((img)[(size_t)(w) * (size_t)(y) + (size_t)(x)])

static float fminf2(float a, float b) { return (b < a) ? b : a; } /* std::min */
static float fmaxf2(float a, float b) { return (a < b) ? b : a; } /* std::max */

/* voxel_carver.cc:102-167 */
void vo_distance_transform_l1(const uint8_t* mask, int w, int h, int rx0, int ry0, int rx1,
                              int ry1, float* d) {
  int x, y;
  for (x = 0; x < w * h; x++) d[x] = 0.0f; /* :104 */
  for (y = ry0; y <= ry1; y++) {           /* :107-114 */
    for (x = rx0; x <= rx1; x++) {
      if (AT(mask, w, x, y) != 255) continue;
      AT(d, w, x, y) = FLT_MAX;
    }
  }
  /* forward :117-140 */
  for (y = ry0 + 1; y <= ry1; y++) {
    float up = AT(d, w, rx0, y - 1);
    if (up < FLT_MAX) AT(d, w, rx0, y) = fminf2(up + 1.0f, AT(d, w, rx0, y));
  }
  for (x = rx0 + 1; x <= rx1; x++) {
    float left = AT(d, w, x - 1, ry0);
    if (left < FLT_MAX) AT(d, w, x, ry0) = fminf2(left + 1.0f, AT(d, w, x, ry0));
  }
  for (y = ry0 + 1; y <= ry1; y++) {
    for (x = rx0 + 1; x <= rx1; x++) {
      float up = AT(d, w, x, y - 1);
      float left = AT(d, w, x - 1, y);
      float m = fminf2(up, left);
      if (m < FLT_MAX) AT(d, w, x, y) = fminf2(m + 1.0f, AT(d, w, x, y));
    }
  }
  /* backward :143-166 */
  for (y = ry1 - 1; ry0 <= y; y--) {
    float down = AT(d, w, rx1, y + 1);
    if (down < FLT_MAX) AT(d, w, rx1, y) = fminf2(down + 1.0f, AT(d, w, rx1, y));
  }
  for (x = rx1 - 1; rx0 <= x; x--) {
    float right = AT(d, w, x + 1, ry1);
    if (right < FLT_MAX) AT(d, w, x, ry1) = fminf2(right + 1.0f, AT(d, w, x, ry1));
  }
  for (y = ry1 - 1; ry0 <= y; y--) {
    for (x = rx1 - 1; rx0 <= x; x--) {
      float down = AT(d, w, x, y + 1);
      float right = AT(d, w, x + 1, y);
      float m = fminf2(down, right);
      if (m < FLT_MAX) AT(d, w, x, y) = fminf2(m + 1.0f, AT(d, w, x, y));
    }
  }
}

/* voxel_carver.cc:169-237 */
void vo_make_sdf(const uint8_t* mask, int w, int h, int rx0, int ry0, int rx1, int ry1,
                 int normalize, int truncate, float band, float* dist) {
  size_t n = (size_t)w * (size_t)h, i;
  int x, y;
  uint8_t* inv = (uint8_t*)malloc(n ? n : 1);
  float* pos = (float*)malloc((n ? n : 1) * sizeof(float));
  vo_distance_transform_l1(mask, w, h, rx0, ry0, rx1, ry1, dist); /* :175 */
  for (y = ry0; y <= ry1; y++)                                     /* :176-182 */
    for (x = rx0; x <= rx1; x++)
      if (AT(dist, w, x, y) > 0) AT(dist, w, x, y) *= -1;
  memcpy(inv, mask, n); /* :184-193 */
  for (y = ry0; y <= ry1; y++)
    for (x = rx0; x <= rx1; x++) AT(inv, w, x, y) = (AT(inv, w, x, y) == 255) ? 0 : 255;
  vo_distance_transform_l1(inv, w, h, rx0, ry0, rx1, ry1, pos); /* :196 */
  for (y = ry0; y <= ry1; y++)                                   /* :197-203 */
    for (x = rx0; x <= rx1; x++)
      if (AT(inv, w, x, y) == 255) AT(dist, w, x, y) = AT(pos, w, x, y);
  if (normalize && n > 0) { /* :205-222, min/max over the WHOLE image */
    float mx = dist[0], mn = dist[0], abs_max;
    for (i = 1; i < n; i++) {
      if (mx < dist[i]) mx = dist[i];
      if (dist[i] < mn) mn = dist[i];
    }
    abs_max = fmaxf2(fabsf(mx), fabsf(mn));
    if (abs_max > FLT_MIN) {
      float norm = 1.0f / abs_max;
      for (y = ry0; y <= ry1; y++)
        for (x = rx0; x <= rx1; x++) AT(dist, w, x, y) *= norm;
    }
  }
  if (truncate) { /* :225-236 */
    for (y = ry0; y <= ry1; y++) {
      for (x = rx0; x <= rx1; x++) {
        float* d = &AT(dist, w, x, y);
        if (-band >= *d) {
          *d = -FLT_MAX;
        } else {
          *d = fminf2(1.0f, *d / band);
        }
      }
    }
  }
  free(inv);
  free(pos);
}

/* Exact squared Euclidean distance to the nearest pixel with seed[i] != 0 inside the ROI, for every
 * ROI pixel (-1 when there is none): 1-D distances along the rows, then the lower envelope of
 * parabolas along each column (Meijster, Roerdink, Hesselink 2000), integers only. */
static void edt_squared(const uint8_t* seed, int w, int rx0, int ry0, int rx1, int ry1,
                        long long* d2) {
  const int rw = rx1 - rx0 + 1, rh = ry1 - ry0 + 1;
  const long long INF = 1LL << 40;
  long long* g = (long long*)malloc((size_t)rw * rh * sizeof(long long));
  int* s = (int*)malloc((size_t)rh * sizeof(int));
  int* t = (int*)malloc((size_t)rh * sizeof(int));
  int x, y;
  for (y = 0; y < rh; y++) { /* nearest seed along the row */
    long long last = -INF;
    for (x = 0; x < rw; x++) {
      if (AT(seed, w, rx0 + x, ry0 + y)) last = x;
      g[(size_t)y * rw + x] = x - last;
    }
    last = INF;
    for (x = rw - 1; x >= 0; x--) {
      if (AT(seed, w, rx0 + x, ry0 + y)) last = x;
      if (last - x < g[(size_t)y * rw + x]) g[(size_t)y * rw + x] = last - x;
    }
  }
  for (x = 0; x < rw; x++) {
    int q = -1, u;
#define G2(i) (g[(size_t)(i) * rw + x] * g[(size_t)(i) * rw + x])
#define F(px, i) (((long long)(px) - (i)) * ((long long)(px) - (i)) + G2(i))
    for (u = 0; u < rh; u++) {
      if (g[(size_t)u * rw + x] >= INF / 2) continue;
      while (q >= 0 && F(t[q], s[q]) > F(t[q], u)) q--;
      if (q < 0) {
        q = 0;
        s[0] = u;
        t[0] = 0;
      } else {
        long long num = (long long)u * u - (long long)s[q] * s[q] + G2(u) - G2(s[q]);
        long long wv = 1 + num / (2 * (u - s[q]));
        if (wv < rh) {
          q++;
          s[q] = u;
          t[q] = (int)wv;
        }
      }
    }
    for (u = rh - 1; u >= 0; u--) {
      d2[(size_t)(ry0 + u) * w + (rx0 + x)] = q >= 0 ? F(u, s[q]) : -1;
      if (q >= 0 && u == t[q]) q--;
    }
#undef F
#undef G2
  }
  free(g);
  free(s);
  free(t);
}

void vo_make_sdf_l2(const uint8_t* mask, int w, int h, int rx0, int ry0, int rx1, int ry1,
                    int normalize, int truncate, float band, float* dist) {
  size_t n = (size_t)w * (size_t)h, i;
  int x, y;
  uint8_t* fg = (uint8_t*)calloc(n ? n : 1, 1);
  uint8_t* bg = (uint8_t*)calloc(n ? n : 1, 1);
  long long* to_bg = (long long*)malloc((n ? n : 1) * sizeof(long long));
  long long* to_fg = (long long*)malloc((n ? n : 1) * sizeof(long long));
  for (i = 0; i < n; i++) {
    fg[i] = mask[i] == 255;
    bg[i] = !fg[i];
    dist[i] = 0.0f;
  }
  edt_squared(bg, w, rx0, ry0, rx1, ry1, to_bg);
  edt_squared(fg, w, rx0, ry0, rx1, ry1, to_fg);
  for (y = ry0; y <= ry1; y++) {
    for (x = rx0; x <= rx1; x++) {
      long long d2 = AT(fg, w, x, y) ? AT(to_bg, w, x, y) : AT(to_fg, w, x, y);
      float e = d2 < 0 ? FLT_MAX : (float)sqrt((double)d2);
      AT(dist, w, x, y) = AT(fg, w, x, y) ? -e : e;
    }
  }
  if (normalize && n > 0) { /* as voxel_carver.cc:205-222 */
    float mx = dist[0], mn = dist[0], abs_max;
    for (i = 1; i < n; i++) {
      if (mx < dist[i]) mx = dist[i];
      if (dist[i] < mn) mn = dist[i];
    }
    abs_max = fmaxf2(fabsf(mx), fabsf(mn));
    if (abs_max > FLT_MIN) {
      float norm = 1.0f / abs_max;
      for (y = ry0; y <= ry1; y++)
        for (x = rx0; x <= rx1; x++) AT(dist, w, x, y) *= norm;
    }
  }
  if (truncate) { /* as voxel_carver.cc:225-236 */
    for (y = ry0; y <= ry1; y++) {
      for (x = rx0; x <= rx1; x++) {
        float* d = &AT(dist, w, x, y);
        if (-band >= *d) {
          *d = -FLT_MAX;
        } else {
          *d = fminf2(1.0f, *d / band);
        }
      }
    }
  }
  free(fg);
  free(bg);
  free(to_bg);
  free(to_fg);
}

/* voxel_carver.cc:239-267 */
void vo_signed_distance_to_color(const float* sdf, int w, int h, float min_negative_d,
                                 float max_positive_d, uint8_t* rgb) {
  size_t i, n = (size_t)w * (size_t)h;
  for (i = 0; i < n; i++) {
    float d = sdf[i];
    if (d > 0) {
      float t = (max_positive_d - d) / max_positive_d;
      t = fminf2(fmaxf2(t, 0.0f), 1.0f);
      rgb[3 * i + 0] = (uint8_t)255;
      rgb[3 * i + 1] = (uint8_t)(255 * t);
      rgb[3 * i + 2] = (uint8_t)(255 * t);
    } else {
      float t = (d - min_negative_d) / (-min_negative_d);
      t = fminf2(fmaxf2(t, 0.0f), 1.0f);
      rgb[3 * i + 0] = (uint8_t)(255 * t);
      rgb[3 * i + 1] = (uint8_t)(255 * t);
      rgb[3 * i + 2] = (uint8_t)255;
    }
  }
}

/* ---- grid ----------------------------------------------------------------------------------- */

/* voxel_carver.cc:375-389 (option checks) and :276-301 (grid checks and sizing) */
int vo_grid_dims(const vo_options* o, int* dims) {
  int i;
  if (o->voxel_max_update_num < 1) return 0;
  if (o->voxel_update_weight < FLT_MIN) return 0;
  if (o->truncation_band < FLT_MIN) return 0;
  if (o->resolution < FLT_MIN) return 0;
  if (o->bb_max[0] <= o->bb_min[0] || o->bb_max[1] <= o->bb_min[1] || o->bb_max[2] <= o->bb_min[2])
    return 0;
  for (i = 0; i < 3; i++) {
    float diff = o->bb_max[i] - o->bb_min[i];
    dims[i] = (int)(diff / o->resolution); /* :294-296 truncation */
  }
  return 1;
}

/* voxel_carver.cc:309-326: pos = diff * (float(k)/float(n)) + bb_min + resolution*0.5f */
void vo_grid_positions(const vo_options* o, float* px, float* py, float* pz) {
  int dims[3], a, k;
  float* out[3];
  float offset = o->resolution * 0.5f;
  out[0] = px;
  out[1] = py;
  out[2] = pz;
  if (!vo_grid_dims(o, dims)) return;
  for (a = 0; a < 3; a++) {
    float diff = o->bb_max[a] - o->bb_min[a];
    for (k = 0; k < dims[a]; k++) {
      float frac = (float)k / (float)dims[a];
      float t = diff * frac;
      t = t + o->bb_min[a];
      out[a][k] = t + offset;
    }
  }
}

void vo_grid_reset(float* sdf, int32_t* update_num, int64_t count) {
  int64_t i;
  for (i = 0; i < count; i++) {
    sdf[i] = -FLT_MAX;
    update_num[i] = 0;
  }
}

/* ---- fusion --------------------------------------------------------------------------------- */

/* voxel_carver.cc:16-38 */
static float interp_nn(float u, float v, const float* img, int w, const int* rmin, const int* rmax) {
  int xi = (int)roundf(u), yi = (int)roundf(v);
  if (xi < rmin[0]) xi = rmin[0];
  if (yi < rmin[1]) yi = rmin[1];
  if (rmax[0] < xi) xi = rmax[0];
  if (rmax[1] < yi) yi = rmax[1];
  return AT(img, w, xi, yi);
}

/* voxel_carver.cc:40-76 */
static float interp_bilinear(float u, float v, const float* img, int w, const int* rmin,
                             const int* rmax) {
  int x0 = (int)floorf(u), y0 = (int)floorf(v);
  int x1 = x0 + 1, y1 = y0 + 1;
  float lu, lv, dist;
  if (x0 < rmin[0]) x0 = rmin[0];
  if (y0 < rmin[1]) y0 = rmin[1];
  if (rmax[0] < x1) x1 = rmax[0];
  if (rmax[1] < y1) y1 = rmax[1];
  lu = u - (float)x0;
  lv = v - (float)y0;
  dist = (1.0f - lu) * (1.0f - lv) * AT(img, w, x0, y0) + lu * (1.0f - lv) * AT(img, w, x1, y0) +
         (1.0f - lu) * lv * AT(img, w, x0, y1) + lu * lv * AT(img, w, x1, y1);
  return dist;
}

/* voxel_carver.cc:415-496 */
void vo_carve_sdf(const vo_options* o, const int* dims, const float* px, const float* py,
                  const float* pz, int z_begin, int z_end, const vo_view* view, const float* img,
                  float* sdf, int32_t* update_num) {
  const int nx = dims[0], ny = dims[1];
  const int w = view->width;
  const float* M = view->w2c;
  const float wgt = o->voxel_update_weight;
  float max_sdf;
  int z;
  {
    size_t n = (size_t)view->width * (size_t)view->height, i; /* :436 */
    max_sdf = img[0];
    for (i = 1; i < n; i++)
      if (max_sdf < img[i]) max_sdf = img[i];
  }
#pragma omp parallel for schedule(dynamic, 1)
  for (z = z_begin; z < z_end; z++) {
    int x, y;
    for (y = 0; y < ny; y++) {
      for (x = 0; x < nx; x++) {
        size_t id = (size_t)z * nx * ny + ((size_t)y * nx + x);
        float pcx, pcy, pcz, u, v, dist;
        int32_t n = update_num[id];
        if (n > o->voxel_max_update_num) continue; /* :447-450 */
        /* :453, Eigen Affine3f * Vector3f: ((m0*x + m1*y) + m2*z) + t */
        pcx = ((M[0] * px[x] + M[1] * py[y]) + M[2] * pz[z]) + M[3];
        pcy = ((M[4] * px[x] + M[5] * py[y]) + M[6] * pz[z]) + M[7];
        pcz = ((M[8] * px[x] + M[9] * py[y]) + M[10] * pz[z]) + M[11];
        if (pcz < 0) continue; /* :456 */
        if (view->model == 0) { /* camera.cc:131-137 */
          u = view->fx / pcz * pcx + view->cx;
          v = view->fy / pcz * pcy + view->cy;
        } else { /* camera.cc:201-205 */
          u = pcx;
          v = pcy;
        }
        dist = -FLT_MAX;
        if (u < view->roi_min[0] || v < view->roi_min[1] || view->roi_max[0] < u ||
            view->roi_max[1] < v) { /* :464-472 */
          if (o->update_outside == 0) {
            continue;
          } else if (o->update_outside == 1) {
            dist = max_sdf;
          }
        } else {
          dist = (o->sdf_interp == 0)
                     ? interp_nn(u, v, img, w, view->roi_min, view->roi_max)
                     : interp_bilinear(u, v, img, w, view->roi_min, view->roi_max);
        }
        if (o->use_truncation && dist < -1.0f) continue; /* :478-480 */
        if (n < 1) { /* :482-486 */
          sdf[id] = dist;
          update_num[id] = n + 1;
          continue;
        }
        if (o->voxel_update == 0) { /* :78-86 */
          if (dist > sdf[id]) {
            sdf[id] = dist;
            update_num[id] = n + 1;
          }
        } else { /* :88-95 */
          const float inv_denom = 1.0f / (wgt * (n + 1));
          sdf[id] = (wgt * n * sdf[id] + wgt * dist) * inv_denom;
          update_num[id] = n + 1;
        }
      }
    }
  }
}

/* ---- ExtractVoxel --------------------------------------------------------------------------- */

/* extract_voxel.cc:258-317; on_surface rule :15-79 */
int64_t vo_extract_voxel_flags(const int* dims, const float* sdf, const int32_t* un,
                               int inside_empty, uint8_t* flags) {
  const int nx = dims[0], ny = dims[1], nz = dims[2];
  const size_t sx = 1, sy = (size_t)nx, sz = (size_t)nx * ny;
  int64_t count = 0;
  int x, y, z;
  for (z = 0; z < nz; z++) {
    for (y = 0; y < ny; y++) {
      for (x = 0; x < nx; x++) {
        size_t id = (size_t)z * sz + (size_t)y * sy + x;
        int emit;
        if (inside_empty) {
          int a;
          const int coord[3] = {x, y, z};
          const size_t stride[3] = {sx, sy, sz};
          emit = 0;
          for (a = 0; a < 3; a++) {
            size_t pid;
            if (coord[a] < 1) continue;
            pid = id - stride[a];
            if (un[id] < 1 || un[pid] < 1) continue;
            if (sdf[id] * sdf[pid] < 0) emit = 1;
            if (fabsf(sdf[id]) < FLT_MIN) emit = 1;
          }
        } else {
          emit = !(sdf[id] > 0 || un[id] < 1); /* :285 */
        }
        flags[id] = (uint8_t)emit;
        count += emit;
      }
    }
  }
  return count;
}

/* mesh.cc:728-780 corner signs (x,y,z) of the 24 cube vertices and the 12 faces */
static const signed char kCubeSign[24][3] = {
    {-1, 1, -1}, {1, 1, -1},  {1, 1, 1},   {-1, 1, 1},   /* 0-3   +y face */
    {-1, -1, -1}, {1, -1, -1}, {1, -1, 1},  {-1, -1, 1}, /* 4-7   -y face */
    {1, 1, -1},  {1, 1, 1},   {1, -1, 1},  {1, -1, -1},  /* 8-11  = v1 v2 v6 v5 */
    {-1, 1, -1}, {-1, 1, 1},  {-1, -1, 1}, {-1, -1, -1}, /* 12-15 = v0 v3 v7 v4 */
    {-1, 1, -1}, {1, 1, -1},  {1, -1, -1}, {-1, -1, -1}, /* 16-19 = v0 v1 v5 v4 */
    {-1, 1, 1},  {1, 1, 1},   {1, -1, 1},  {-1, -1, 1}}; /* 20-23 = v3 v2 v6 v7 */
static const int kCubeFace[12][3] = {{0, 2, 1},    {0, 3, 2},    {4, 5, 6},    {4, 6, 7},
                                     {8, 9, 10},   {8, 10, 11},  {12, 14, 13}, {12, 15, 14},
                                     {16, 17, 18}, {16, 18, 19}, {20, 22, 21}, {20, 23, 22}};

void vo_extract_voxel_mesh(const int* dims, const float* px, const float* py, const float* pz,
                           float resolution, const uint8_t* flags, float* vertices,
                           int32_t* faces) {
  const int nx = dims[0], ny = dims[1], nz = dims[2];
  const float h = resolution / 2; /* mesh.cc:736-738 */
  int64_t k = 0;
  int x, y, z, i;
  for (z = 0; z < nz; z++) {
    for (y = 0; y < ny; y++) {
      for (x = 0; x < nx; x++) {
        size_t id = ((size_t)z * ny + y) * nx + x;
        if (!flags[id]) continue;
        for (i = 0; i < 24; i++) {
          /* cube->Translate(pos): v + t, mesh.cc:161-166 */
          vertices[(k * 24 + i) * 3 + 0] = (kCubeSign[i][0] < 0 ? -h : h) + px[x];
          vertices[(k * 24 + i) * 3 + 1] = (kCubeSign[i][1] < 0 ? -h : h) + py[y];
          vertices[(k * 24 + i) * 3 + 2] = (kCubeSign[i][2] < 0 ? -h : h) + pz[z];
        }
        for (i = 0; i < 12; i++) {
          faces[(k * 12 + i) * 3 + 0] = (int32_t)(kCubeFace[i][0] + 24 * k);
          faces[(k * 12 + i) * 3 + 1] = (int32_t)(kCubeFace[i][1] + 24 * k);
          faces[(k * 12 + i) * 3 + 2] = (int32_t)(kCubeFace[i][2] + 24 * k);
        }
        k++;
      }
    }
  }
}

/* ---- marching cubes ------------------------------------------------------------------------- */

/* corner k -> offset from the cube's max corner (x,y,z), marching_cubes.cc:93-101 */
static const int kCorner[8][3] = {{-1, -1, -1}, {0, -1, -1}, {0, 0, -1}, {-1, 0, -1},
                                  {-1, -1, 0},  {0, -1, 0},  {0, 0, 0},  {-1, 0, 0}};
/* edge e -> (interpolate from corner, to corner), marching_cubes.cc:138-197 */
static const int kEdgeCorner[12][2] = {{0, 1}, {1, 2}, {2, 3}, {3, 0}, {4, 5}, {5, 6},
                                       {6, 7}, {7, 4}, {0, 4}, {1, 5}, {2, 6}, {3, 7}};

/* An edge is flagged iff its two corners lie on different sides (Bourke's edgeTable) */
static int edge_mask_of(int cube_index) {
  int e, m = 0;
  for (e = 0; e < 12; e++) {
    int a = (cube_index >> kEdgeCorner[e][0]) & 1, b = (cube_index >> kEdgeCorner[e][1]) & 1;
    if (a != b) m |= 1 << e;
  }
  return m;
}

void vo_mc_tables(int* edge_table, int* tri_table) {
  int c, i;
  for (c = 0; c < 256; c++) {
    edge_table[c] = edge_mask_of(c);
    for (i = 0; i < 16; i++) {
      int nib = (int)((kTriPacked[c] >> (4 * i)) & 0xF);
      tri_table[c * 16 + i] = (nib == 0xF) ? -1 : nib;
    }
  }
}

/* marching_cubes.cc:25-47 */
static void vertex_interp(double iso, const float* p1, const float* p2, double v1, double v2,
                          float* p) {
  double mu;
  int a;
  if (fabs(iso - v1) < 0.00001) {
    memcpy(p, p1, 3 * sizeof(float));
    return;
  }
  if (fabs(iso - v2) < 0.00001) {
    memcpy(p, p2, 3 * sizeof(float));
    return;
  }
  if (fabs(v1 - v2) < 0.00001) {
    memcpy(p, p1, 3 * sizeof(float));
    return;
  }
  mu = (iso - v1) / (v2 - v1);
  for (a = 0; a < 3; a++) p[a] = (float)(p1[a] + mu * ((double)p2[a] - (double)p1[a]));
}

/* std::map<pair<int,int>,int> stand-in: open addressing on the (lo id, hi id) key. Only find and
 * insert are used by the reference (marching_cubes.cc:204-215), so ordering is irrelevant. */
typedef struct {
  uint64_t* keys;
  int32_t* vals;
  size_t cap, used;
} edge_map;

static size_t em_slot(const edge_map* m, uint64_t key) {
  uint64_t hsh = key * 0x9E3779B97F4A7C15ULL;
  size_t i = (size_t)(hsh >> 17) & (m->cap - 1);
  while (m->keys[i] != UINT64_MAX && m->keys[i] != key) i = (i + 1) & (m->cap - 1);
  return i;
}

static void em_grow(edge_map* m) {
  edge_map n;
  size_t i;
  n.cap = m->cap ? m->cap * 2 : 1024;
  n.used = m->used;
  n.keys = (uint64_t*)malloc(n.cap * sizeof(uint64_t));
  n.vals = (int32_t*)malloc(n.cap * sizeof(int32_t));
  memset(n.keys, 0xFF, n.cap * sizeof(uint64_t));
  for (i = 0; i < m->cap; i++) {
    if (m->keys[i] != UINT64_MAX) {
      size_t s = em_slot(&n, m->keys[i]);
      n.keys[s] = m->keys[i];
      n.vals[s] = m->vals[i];
    }
  }
  free(m->keys);
  free(m->vals);
  *m = n;
}

/* marching_cubes.cc:63-228 */
void vo_marching_cubes(const int* dims, const float* px, const float* py, const float* pz,
                       const float* sdf, const int32_t* un, double iso, int linear_interp,
                       float** vertices_out, int64_t* n_vertices, int32_t** faces_out,
                       int64_t* n_faces) {
  const int nx = dims[0], ny = dims[1], nz = dims[2];
  const float kVal = -FLT_MAX;
  size_t vcap = 1024, fcap = 1024, nv = 0, nf = 0;
  float* verts = (float*)malloc(vcap * 3 * sizeof(float));
  int32_t* faces = (int32_t*)malloc(fcap * 3 * sizeof(int32_t));
  edge_map map = {0, 0, 0, 0};
  int x, y, z, k, e, i, j;
  em_grow(&map);
  for (z = 1; z < nz; z++) {
    for (y = 1; y < ny; y++) {
      for (x = 1; x < nx; x++) {
        size_t id[8];
        float val[8], pos[8][3], vert_list[12][3];
        uint64_t key_list[12];
        int cube_index = 0, em, skip = 0;
        uint64_t tri;
        if (un[((size_t)z * ny + y) * nx + x] < 1) continue; /* :88-90 */
        for (k = 0; k < 8; k++) {
          int cx = x + kCorner[k][0], cy = y + kCorner[k][1], cz = z + kCorner[k][2];
          id[k] = ((size_t)cz * ny + cy) * nx + cx;
          val[k] = sdf[id[k]];
          pos[k][0] = px[cx];
          pos[k][1] = py[cy];
          pos[k][2] = pz[cz];
          if (val[k] == kVal) skip = 1; /* :103-112 */
        }
        if (skip) continue;
        for (k = 0; k < 8; k++)
          if ((double)val[k] < iso) cube_index |= 1 << k; /* :121-128 */
        em = edge_mask_of(cube_index);
        if (em == 0) continue; /* :131-133 */
        for (e = 0; e < 12; e++) { /* :138-197 */
          int a, b;
          uint64_t lo, hi;
          if (!(em & (1 << e))) continue;
          a = kEdgeCorner[e][0];
          b = kEdgeCorner[e][1];
          if (linear_interp) {
            vertex_interp(iso, pos[a], pos[b], (double)val[a], (double)val[b], vert_list[e]);
          } else {
            memcpy(vert_list[e], pos[a], 3 * sizeof(float)); /* :55 */
          }
          lo = id[a] < id[b] ? id[a] : id[b];
          hi = id[a] < id[b] ? id[b] : id[a];
          key_list[e] = (lo << 32) | hi;
        }
        tri = kTriPacked[cube_index];
        for (i = 0; ((tri >> (4 * i)) & 0xF) != 0xF; i += 3) { /* :199-218 */
          int32_t face[3];
          for (j = 0; j < 3; j++) {
            int edge = (int)((tri >> (4 * (i + (2 - j)))) & 0xF);
            size_t s = em_slot(&map, key_list[edge]);
            if (map.keys[s] == UINT64_MAX) {
              if (nv + 1 > vcap) {
                vcap *= 2;
                verts = (float*)realloc(verts, vcap * 3 * sizeof(float));
              }
              face[j] = (int32_t)nv;
              memcpy(&verts[nv * 3], vert_list[edge], 3 * sizeof(float));
              nv++;
              map.keys[s] = key_list[edge];
              map.vals[s] = face[j];
              map.used++;
              if (map.used * 2 > map.cap) em_grow(&map);
            } else {
              face[j] = map.vals[s];
            }
          }
          if (nf + 1 > fcap) {
            fcap *= 2;
            faces = (int32_t*)realloc(faces, fcap * 3 * sizeof(int32_t));
          }
          memcpy(&faces[nf * 3], face, 3 * sizeof(int32_t));
          nf++;
          if (i + 3 >= 16) break;
        }
      }
    }
  }
  free(map.keys);
  free(map.vals);
  *vertices_out = verts;
  *n_vertices = (int64_t)nv;
  *faces_out = faces;
  *n_faces = (int64_t)nf;
}

void vo_free(void* p) { free(p); }

/* ---- Mesh::WritePly, mesh.cc:583-631 (no vertex colours) ------------------------------------- */
int64_t vo_ply_text(const float* vertices, int64_t n_vertices, const int32_t* faces, int64_t n_faces,
                    char* text, int64_t capacity) {
  char line[320];
  int64_t n = 0, i;
  int k;
  /* header, mesh.cc:596-612 */
  k = snprintf(line, sizeof(line),
               "ply\nformat ascii 1.0\nelement vertex %lld\nproperty float x\nproperty float y\n"
               "property float z\nelement face %lld\nproperty list uchar int vertex_indices\n"
               "end_header\n",
               (long long)n_vertices, (long long)n_faces);
  if (text && n + k <= capacity) memcpy(text + n, line, (size_t)k);
  n += k;
  for (i = 0; i < n_vertices; i++) { /* mesh.cc:614-624 */
    k = snprintf(line, sizeof(line), "%g %g %g \n", (double)vertices[3 * i],
                 (double)vertices[3 * i + 1], (double)vertices[3 * i + 2]);
    if (text && n + k <= capacity) memcpy(text + n, line, (size_t)k);
    n += k;
  }
  for (i = 0; i < n_faces; i++) { /* mesh.cc:626-629 */
    k = snprintf(line, sizeof(line), "3 %d %d %d \n", faces[3 * i], faces[3 * i + 1],
                 faces[3 * i + 2]);
    if (text && n + k <= capacity) memcpy(text + n, line, (size_t)k);
    n += k;
  }
  return n;
}
