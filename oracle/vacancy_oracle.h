/* TEST INFRASTRUCTURE ONLY - the parity oracle. Never linked into, imported by or called from
 * the product path (vacancy_b200/); only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference leg may use it, and only as the checker.
 *
 * Plain-C restatement of the carving hot path of unclearness/vacancy (reference commit 4af8aecf).
 * Every function cites the reference file:line it follows; arithmetic is float32 / float64 in the
 * reference's exact operation order, compiled with -ffp-contract=off (no FMA), so results are meant
 * to be BIT-IDENTICAL to the reference.  Pinned by tests/test_oracle_*.py against the compiled
 * reference itself (oracle/_ref, built from /root/reference) and against the fixtures under
 * tests/golden/ that were generated from it (tests/golden/make_golden.py).
 *
 * Grid state is SoA: sdf[z*nx*ny + y*nx + x] (x fastest, voxel_carver.cc:350), update_num[] same.
 */
#ifndef VACANCY_ORACLE_H_
#define VACANCY_ORACLE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vo_options { /* voxel_carver.h:43-60 */
  float bb_max[3];
  float bb_min[3];
  float resolution;
  int32_t sdf_minmax_normalize;
  int32_t voxel_update;   /* 0 kMax, 1 kWeightedAverage (voxel_carver.h:20-24) */
  int32_t sdf_interp;     /* 0 kNn, 1 kBilinear (voxel_carver.h:27-30) */
  int32_t update_outside; /* 0 kNone, 1 kMax (voxel_carver.h:33-37) */
  int32_t voxel_max_update_num;
  float voxel_update_weight;
  int32_t use_truncation;
  float truncation_band;
} vo_options;

typedef struct vo_view {
  float w2c[12];   /* row-major 3x4 = camera.w2c().cast<float>() (voxel_carver.cc:438) */
  int32_t model;   /* 0 pinhole (camera.cc:131-137), 1 ortho (camera.cc:201-205) */
  float fx, fy, cx, cy;
  int32_t width, height;
  int32_t roi_min[2], roi_max[2]; /* inclusive */
} vo_view;

int vo_set_threads(int n); /* OpenMP threads of the fusion loop; n < 0 = all cores, 0 = report only */
float vo_invalid_sdf(void); /* InvalidSdf::kVal voxel_carver.cc:100 */

/* camera.cc:39-42 + Eigen 3.4.0 Transform::inverse(Affine) + cast<float> */
void vo_w2c_from_c2w(const double* c2w16, float* w2c12);

/* voxel_carver.cc:102-167 */
void vo_distance_transform_l1(const uint8_t* mask, int w, int h, int rx0, int ry0, int rx1,
                              int ry1, float* out);
/* voxel_carver.cc:169-237 */
void vo_make_sdf(const uint8_t* mask, int w, int h, int rx0, int ry0, int rx1, int ry1,
                 int normalize, int truncate, float band, float* out);

/* NOT in the reference: the same signed field with the exact Euclidean metric (the product's opt-in
 * VCB_METRIC_L2). Checker for that mode only; pinned against scipy.ndimage.distance_transform_edt and
 * a brute-force search in tests/test_oracle_pinned.py, not against the reference (which is L1).
 * Distance = (float)sqrt((double)squared integer distance); sign, FLT_MAX convention, normalisation
 * and truncation as in vo_make_sdf. */
void vo_make_sdf_l2(const uint8_t* mask, int w, int h, int rx0, int ry0, int rx1, int ry1,
                    int normalize, int truncate, float band, float* out);

/* SignedDistance2Color, voxel_carver.cc:239-267; rgb = 3 bytes per pixel */
void vo_signed_distance_to_color(const float* sdf, int w, int h, float min_negative_d,
                                 float max_positive_d, uint8_t* rgb);

/* VoxelCarver::Init + VoxelGrid::Init validation and sizing, voxel_carver.cc:276-301,375-389.
 * Returns 1 on success (dims filled), 0 when the reference would return false. */
int vo_grid_dims(const vo_options* opt, int* dims3);
/* voxel centres, voxel_carver.cc:315-326 */
void vo_grid_positions(const vo_options* opt, float* px, float* py, float* pz);
/* sdf = kVal, update_num = 0, voxel_carver.cc:342 and voxel_carver.h:67-68 */
void vo_grid_reset(float* sdf, int32_t* update_num, int64_t count);

/* Carve(camera, roi_min, roi_max, sdf) main loop, voxel_carver.cc:415-496, applied to the
 * z planes [z_begin, z_end) of the grid (the whole grid is 0..nz). */
void vo_carve_sdf(const vo_options* opt, const int* dims3, const float* px, const float* py,
                  const float* pz, int z_begin, int z_end, const vo_view* view,
                  const float* sdf_image, float* sdf, int32_t* update_num);

/* ExtractVoxel, extract_voxel.cc:258-317 (+ UpdateOnSurface :15-79 when inside_empty).
 * flags[] (one byte per voxel) receives 1 for every emitted voxel; returns the count. */
int64_t vo_extract_voxel_flags(const int* dims3, const float* sdf, const int32_t* update_num,
                               int inside_empty, uint8_t* flags);
/* 24 vertices + 12 faces per flagged voxel in raster order; cube geometry mesh.cc:728-816.
 * Vertices are pos + corner (the reference's translate/untranslate drift is not modelled). */
void vo_extract_voxel_mesh(const int* dims3, const float* px, const float* py, const float* pz,
                           float resolution, const uint8_t* flags, float* vertices,
                           int32_t* faces);

/* MarchingCubes, marching_cubes.cc:63-228. Allocates *vertices (3 floats each) and *faces
 * (3 int32 each) with malloc; free with vo_free. */
void vo_marching_cubes(const int* dims3, const float* px, const float* py, const float* pz,
                       const float* sdf, const int32_t* update_num, double iso_level,
                       int linear_interp, float** vertices, int64_t* n_vertices,
                       int32_t** faces, int64_t* n_faces);
void vo_free(void* p);

/* Mesh::WritePly, mesh.cc:583-631, for a mesh without vertex colours: writes the file's bytes into
 * text (capacity bytes) and returns their number (the needed size when text is NULL or too small).
 * `ofs << float` is printf("%g"), `<< int` printf("%d"), std::endl a newline. */
int64_t vo_ply_text(const float* vertices, int64_t n_vertices, const int32_t* faces, int64_t n_faces,
                    char* text, int64_t capacity);

/* tables: packed triangle table and the derived edge table (marching_cubes_lut.cc:15-298) */
void vo_mc_tables(int* edge_table256, int* tri_table256x16);

#ifdef __cplusplus
}
#endif
#endif /* VACANCY_ORACLE_H_ */
