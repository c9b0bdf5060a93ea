/*
 * vacancy_b200 - C ABI of the B200-native voxel-carving hot path.
 *
 * This is the drop-in boundary for the one path this library accelerates: per-view 2D silhouette
 * signed-distance field -> multi-view fusion into the voxel grid -> ExtractVoxel / marching cubes.
 * The reference (unclearness/vacancy) has no FFI of its own: its boundary is the C++ class
 * vacancy::VoxelCarver (include/vacancy/voxel_carver.h:95-128).  Every entry point below names the
 * reference interface it replaces; vacancy_b200/shim/voxel_carver_shim.cc implements that C++ class
 * on top of these calls.
 *
 * Conventions
 *  - plain C, no CUDA or torch types in the signatures; device pointers are passed as void* / typed
 *    pointers and are documented as such; streams are passed as void* (a cudaStream_t).
 *  - every call returns VCB_OK (0) or a negative vcb_status; vcb_last_error() gives the text.
 *    There is no CPU fallback: without a usable CUDA device vcb_create() fails with VCB_ERR_CUDA.
 *  - images are row-major, x fastest: pixel (x, y) at [y * width + x] (image.h:57-74).
 *  - the grid is row-major with x fastest: voxel (x, y, z) at [z*nx*ny + y*nx + x]
 *    (voxel_carver.cc:347-351); state is SoA: float sdf[] and int32 update_num[].
 *  - all work is ordered on one CUDA stream per carver (vcb_set_stream); calls that take or
 *    return HOST buffers are complete on return, calls on DEVICE buffers are stream-ordered.
 *  - a carver is not thread-safe (same as the reference); distinct carvers are independent.
 */
#ifndef VACANCY_B200_H_
#define VACANCY_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Bumped whenever a struct below changes size or layout; compare with vcb_abi_version() before
 * the first call when the library and the host code are built separately. */
#define VCB_ABI_VERSION 2

typedef enum vcb_status {
  VCB_OK = 0,
  VCB_ERR_INVALID = -1,   /* bad argument / option the reference's Init() would reject */
  VCB_ERR_CUDA = -2,      /* CUDA runtime error (text in vcb_last_error) */
  VCB_ERR_STATE = -3,     /* call out of order (e.g. fill before count) */
  VCB_ERR_UNSUPPORTED = -4
} vcb_status;

/* VoxelUpdate / SdfInterpolation / UpdateOutsideImage, voxel_carver.h:20-37 (same values) */
enum { VCB_UPDATE_MAX = 0, VCB_UPDATE_WEIGHTED_AVERAGE = 1 };
enum { VCB_INTERP_NN = 0, VCB_INTERP_BILINEAR = 1 };
enum { VCB_OUTSIDE_NONE = 0, VCB_OUTSIDE_MAX = 1 };
/* Metric of the 2D silhouette distance. L1 (city block) is what the reference computes
 * (DistanceTransformL1, voxel_carver.cc:102-167) and the only mode with reference parity; L2 is the
 * exact Euclidean distance transform, an addition of this library (everything downstream - sign
 * convention, min/max normalisation, truncation, fusion - is unchanged). */
enum { VCB_METRIC_L1 = 0, VCB_METRIC_L2 = 1 };
/* concrete camera models, camera.h:66 (PinholeCamera) and camera.h:116 (OrthoCamera) */
enum { VCB_CAMERA_PINHOLE = 0, VCB_CAMERA_ORTHO = 1 };

/* VoxelCarverOption + VoxelUpdateOption, voxel_carver.h:43-60, field for field, followed by the
 * placement fields this library adds.  Zero-initialise, then call vcb_options_default(). */
typedef struct vcb_options {
  float bb_max[3];
  float bb_min[3];
  float resolution;             /* default 0.1f */
  int32_t sdf_minmax_normalize; /* default 1 */
  int32_t voxel_update;         /* default VCB_UPDATE_MAX */
  int32_t sdf_interp;           /* default VCB_INTERP_BILINEAR */
  int32_t update_outside;       /* default VCB_OUTSIDE_NONE */
  int32_t voxel_max_update_num; /* default 255 */
  float voxel_update_weight;    /* default 1.0f */
  int32_t use_truncation;       /* default 0 */
  float truncation_band;        /* default 0.1f */
  /* --- additions --- */
  int32_t device;  /* CUDA device ordinal (default 0) */
  int32_t z_begin; /* this carver owns the z planes [z_begin, z_end) of the global grid; */
  int32_t z_end;   /* z_begin == z_end == 0 selects the whole grid (single GPU) */
  int32_t sdf_metric; /* VCB_METRIC_L1 (default, the reference's transform) or VCB_METRIC_L2 */
} vcb_options;

/* One view: camera.w2c().cast<float>() (voxel_carver.cc:438) as a row-major 3x4, the projection
 * model with its intrinsics (camera.cc:131-137 / 201-205), the image size and the inclusive ROI
 * (voxel_carver.cc:394-413). */
typedef struct vcb_view {
  float w2c[12];
  int32_t model;
  float fx, fy, cx, cy;
  int32_t width, height;
  int32_t roi_min[2];
  int32_t roi_max[2];
} vcb_view;

typedef struct vcb_carver vcb_carver;

/* Text of the last error on this carver; with NULL, of the last failed vcb_create / free call
 * of the calling thread. Never NULL. */
const char* vcb_last_error(const vcb_carver* c);

/* VCB_ABI_VERSION the library was built with */
int vcb_abi_version(void);

/* InvalidSdf::kVal, voxel_carver.cc:100 */
float vcb_invalid_sdf(void);

/* Fills the reference's defaults (voxel_carver.h:43-60). */
void vcb_options_default(vcb_options* opt);

/* VoxelCarver::set_option + Init, voxel_carver.cc:371-392 and VoxelGrid::Init :276-345.
 * Rejects what Init() rejects (max_update_num < 1, weight / band / resolution not positive,
 * empty bounding box) with VCB_ERR_INVALID.  Unlike the reference, grids above 2^31 voxels are
 * accepted (64-bit indices); a single carver's slab must still hold < 2^31 * 32 voxels. */
int vcb_create(const vcb_options* opt, vcb_carver** out);
/* Init() called again on a live carver (voxel_carver.cc:390-391 re-creates the grid): every voxel
 * back to sdf = kVal, update_num = 0, queued views dropped; buffers are kept. Stream-ordered. */
int vcb_reset(vcb_carver* c);
/* VoxelCarver::~VoxelCarver, voxel_carver.cc:369 */
void vcb_destroy(vcb_carver* c);

/* VoxelGrid::voxel_num, voxel_carver.cc:347: global grid size; slab = owned z range */
int vcb_grid_dims(const vcb_carver* c, int32_t dims3[3]);
int vcb_slab_range(const vcb_carver* c, int32_t* z_begin, int32_t* z_end);
/* Voxel::pos along each axis (voxel_carver.cc:315-326), global tables of nx, ny, nz floats (host) */
int vcb_grid_positions(const vcb_carver* c, float* px, float* py, float* pz);

/* All later work of this carver is issued on `cuda_stream` (a cudaStream_t; NULL = the legacy
 * default stream). Synchronises the previous stream first. */
int vcb_set_stream(vcb_carver* c, void* cuda_stream);
/* Launches any queued fusion work and waits for the carver's stream. */
int vcb_synchronize(vcb_carver* c);
/* Launches the fusion of every view queued so far (views are queued by the carve calls and fused
 * in call order, up to 64 per kernel launch); does not wait. Extract / download calls flush
 * implicitly, so this is only needed for timing. */
int vcb_flush(vcb_carver* c);

/* Carve(camera, silhouette, roi_min, roi_max, sdf*) for n views, voxel_carver.cc:394-413:
 * builds each view's 2D SDF from its mask (MakeSignedDistanceField :169-237 with the carver's
 * normalize / truncation options) and fuses it (:415-496).  masks[i]: HOST uint8 image of
 * views[i].width x height.  sdf_out: NULL, or n HOST float pointers (entries may be NULL) that
 * receive each view's 2D SDF like the reference's `Image1f* sdf` out-parameter. */
int vcb_carve_masks(vcb_carver* c, int32_t n_views, const vcb_view* views,
                    const uint8_t* const* masks, float* const* sdf_out);
/* Same with masks already in DEVICE memory (stream-ordered; masks must stay valid until the
 * carver's stream has passed this call). sdf_out_dev: NULL or n DEVICE float pointers. */
int vcb_carve_masks_device(vcb_carver* c, int32_t n_views, const vcb_view* views,
                           const uint8_t* const* masks_dev, float* const* sdf_out_dev);
/* Carve(camera, roi_min, roi_max, sdf) for n views, voxel_carver.cc:415-496; HOST float SDFs */
int vcb_carve_sdfs(vcb_carver* c, int32_t n_views, const vcb_view* views,
                   const float* const* sdfs);
/* Same with DEVICE float SDF images, fused in place (no copy): flushes earlier queued views,
 * then launches these; the images must stay valid until the stream has passed this call. */
int vcb_carve_sdfs_device(vcb_carver* c, int32_t n_views, const vcb_view* views,
                          const float* const* sdfs_dev);

/* Image1b::Load (image.h:76-101) + Carve for a batch of silhouette FILES: the files are decoded on
 * host threads (n_threads <= 0: one per hardware thread) into page-locked memory, one fusion launch
 * worth at a time, and carved as by vcb_carve_masks - decoding the next files overlaps the GPU
 * work on the previous ones.  Decoded here: 8-bit greyscale non-interlaced PNG (the reference's
 * data/mask_*.png; same pixels as its stb loader gives) and binary PGM (P5, maxval 255); other
 * files are refused with VCB_ERR_UNSUPPORTED.  views[i].width / height must be the image's size.
 * vcb_load_mask_file decodes one file into `pixels` (capacity bytes; NULL: only report the size). */
int vcb_carve_mask_files(vcb_carver* c, int32_t n_views, const vcb_view* views,
                         const char* const* paths, int32_t n_threads);
int vcb_load_mask_file(const char* path, uint8_t* pixels, size_t capacity, int32_t* width,
                       int32_t* height);

/* Free functions of voxel_carver.h:120-126 on HOST buffers, run on CUDA device `device` (the
 * caller's current device is restored before they return; they allocate and free their device
 * buffers per call - hosts that transform many images should queue them on a carver instead).
 * DistanceTransformL1 (voxel_carver.cc:102-167) and MakeSignedDistanceField (:169-237). */
int vcb_distance_transform_l1(int32_t device, const uint8_t* mask, int32_t width, int32_t height,
                              const int32_t roi_min[2], const int32_t roi_max[2], float* dist);
int vcb_make_signed_distance_field(int32_t device, const uint8_t* mask, int32_t width,
                                   int32_t height, const int32_t roi_min[2],
                                   const int32_t roi_max[2], int32_t minmax_normalize,
                                   int32_t use_truncation, float truncation_band, float* dist);

/* No counterpart in the reference: MakeSignedDistanceField with the exact Euclidean metric. For a
 * foreground (== 255) pixel the value is minus the distance between pixel centres to the nearest
 * other pixel of the ROI, for any other pixel plus the distance to the nearest foreground pixel;
 * distances are sqrt(integer) evaluated in float64 and rounded to float32; -+FLT_MAX when the ROI
 * holds no pixel of the other class. Normalisation and truncation as in voxel_carver.cc:205-236. */
int vcb_make_signed_distance_field_l2(int32_t device, const uint8_t* mask, int32_t width,
                                      int32_t height, const int32_t roi_min[2],
                                      const int32_t roi_max[2], int32_t minmax_normalize,
                                      int32_t use_truncation, float truncation_band, float* dist);

/* SignedDistance2Color (voxel_carver.h:127-128, voxel_carver.cc:239-267): the debugging colour map of
 * a 2D SDF, HOST float image in, HOST RGB image (3 bytes per pixel) out; positive distances fade
 * from white to red, the others from white to blue. Requires min_negative_d < 0 < max_positive_d. */
int vcb_signed_distance_to_color(int32_t device, const float* sdf, int32_t width, int32_t height,
                                 float min_negative_d, float max_positive_d, uint8_t* rgb);

/* No counterpart in the reference: MakeSignedDistanceField for callers that must not allocate.
 * Some hosts cannot let a library call cudaMalloc once they are running: an application whose
 * own arena has claimed the GPU's memory (a driver allocation may then fail, or fragment what is
 * left), or an acquisition loop that cannot afford the device-wide synchronisation an allocation
 * implies.  This variant therefore takes the mask and the result as DEVICE pointers supplied by the
 * caller, works on `cuda_stream` (a cudaStream_t created on `device`, NULL = that device's default
 * stream) without waiting for it, and keeps its intermediate storage (two int16 images, the
 * foreground bit-plane, the per-distance value table and two result cells) in a fixed array inside
 * the library.  The caller's current device is the same after the call as before it.
 *   - limit: W32 * height <= 2048 * 2048 pixels, W32 = width rounded up to a multiple of 32;
 *     larger images are refused with VCB_ERR_INVALID (use vcb_make_signed_distance_field);
 *   - cost: every GPU that loads the library holds that array (16.75 MiB) until the process exits;
 *   - the array is shared by all calls and keeps the previous call's contents; each call
 *     re-initialises the cells it accumulates into, but two calls whose kernels overlap on the GPU
 *     (different streams, or different host threads) race on it - serialise them;
 *   - the first call on each device looks up the array's device address (a host-side set-up step);
 *     make that first call before starting any stream capture.
 * The result equals vcb_make_signed_distance_field bit for bit. It never calls cudaMalloc,
 * cudaFree or any allocating routine. */
int vcb_make_signed_distance_field_noalloc(int32_t device, const uint8_t* mask_dev, int32_t width,
                                           int32_t height, const int32_t roi_min[2],
                                           const int32_t roi_max[2], int32_t minmax_normalize,
                                           int32_t use_truncation, float truncation_band,
                                           float* dist_dev, void* cuda_stream);

/* ExtractIsoSurface(mesh, iso_level, linear_interp), voxel_carver.cc:540-543 ->
 * MarchingCubes marching_cubes.cc:63-228, two-phase: _count runs the extraction on the device and
 * reports the sizes, _fill copies the mesh into HOST buffers (3 floats per vertex, 3 int32 per
 * face, same vertex numbering and face order as the reference). */
int vcb_extract_iso_count(vcb_carver* c, double iso_level, int32_t linear_interp,
                          int64_t* n_vertices, int64_t* n_faces);
int vcb_extract_iso_fill(vcb_carver* c, float* vertices, int32_t* faces);
/* DEVICE pointers to the mesh left by the last vcb_extract_iso_count (owned by the carver, valid
 * until its next extract call or destroy). */
int vcb_extract_iso_device(vcb_carver* c, const float** vertices_dev, const int32_t** faces_dev);

/* ASCII PLY text of a mesh, byte for byte what Mesh::WritePly (mesh.cc:583-631) writes for a mesh
 * without vertex colours: the header (:596-612), one line "x y z \n" per vertex (:614-624, floats
 * as `ofs << float`, i.e. "%g" with 6 significant digits) and one line "3 i j k \n" per face
 * (:626-629).  The lines are formatted on the GPU with exact integer arithmetic.
 *   vcb_ply_header     host only; with text == NULL only the size is returned.
 *   vcb_ply_body_size / _fill   the mesh left on the device by the last vcb_extract_iso_count (or
 *       the phased calls): sizes of the vertex lines and of the face lines, then both blocks into
 *       caller buffers of those sizes.  A file is header + vertex block + face block; with z slabs
 *       it is the header for the global counts, every slab's vertex block in slab order, then every
 *       slab's face block (ids are already global).
 *   vcb_ply_format     the whole file from HOST arrays (any mesh); with text == NULL only the size
 *       is returned. */
int vcb_ply_header(int64_t n_vertices, int64_t n_faces, char* text, size_t capacity, size_t* n_bytes);
int vcb_ply_body_size(vcb_carver* c, size_t* vertex_bytes, size_t* face_bytes);
int vcb_ply_body_fill(vcb_carver* c, char* vertex_text, char* face_text);
int vcb_ply_format(int32_t device, const float* vertices, int64_t n_vertices, const int32_t* faces,
                   int64_t n_faces, char* text, size_t capacity, size_t* n_bytes);

/* The same extraction in phases, for a grid sharded into z slabs across GPUs (no reference
 * counterpart: the reference is single-process; numbering still follows marching_cubes.cc:199-218
 * over the WHOLE grid).  Per slab, in this order:
 *   1. fill the halo: the 2 sdf planes below the slab (vcb_halo_device) with the lower slab's top
 *      planes (vcb_grid_device + offset);                                 [neighbour exchange]
 *   2. _begin: classify + count -> this slab's vertex / face counts;
 *   3. all slabs exchange counts; vertex_offset = sum of the lower slabs' vertex counts;
 *   4. _vertices(vertex_offset): emits this slab's vertices, ids become global;
 *   5. _export_top(ids): ids (or -1) of the vertices on the x- and y-edges of the slab's top plane,
 *      2*ny*nx int32 (DEVICE), sent to the slab above, which calls _import_bottom(ids);
 *   6. _faces: emits this slab's faces with global vertex ids.
 * vcb_extract_iso_count == _begin + _vertices(0) + _faces. */
int vcb_extract_iso_begin(vcb_carver* c, double iso_level, int32_t linear_interp,
                          int64_t* n_vertices, int64_t* n_faces);
int vcb_extract_iso_vertices(vcb_carver* c, int64_t vertex_offset);
int vcb_extract_iso_export_top(vcb_carver* c, int32_t* ids_dev);
int vcb_extract_iso_import_bottom(vcb_carver* c, const int32_t* ids_dev);
int vcb_extract_iso_faces(vcb_carver* c);
/* DEVICE pointer to the halo: n_planes (0..2) sdf planes stored directly below the slab
 * (planes z_begin - n_planes .. z_begin - 1), contiguous with the slab's own planes. */
int vcb_halo_device(vcb_carver* c, float** halo_dev, int32_t* n_planes);

/* ExtractVoxel(mesh, inside_empty), voxel_carver.cc:530-538 -> extract_voxel.cc:258-317.
 * _count classifies the voxels and returns how many cubes will be emitted (24 vertices and 12
 * faces each); _fill writes them to HOST buffers in the reference's raster order;
 * _occupancy writes one byte per owned voxel (1 = emitted) to a HOST buffer instead.
 * On a z slab above the first, inside_empty looks at the plane below the slab: fill the halo
 * (vcb_halo_device) from the lower slab first, as for marching cubes (vcb_sharded_* does). */
int vcb_extract_voxel_count(vcb_carver* c, int32_t inside_empty, int64_t* n_voxels);
int vcb_extract_voxel_fill(vcb_carver* c, float* vertices, int32_t* faces);
int vcb_extract_voxel_occupancy(vcb_carver* c, uint8_t* flags);

/* Grid state of the owned slab to / from HOST buffers (the reference keeps the grid private,
 * voxel_carver.h:97; tests and checkpoints use these). Either pointer may be NULL.  Uploading both
 * arrays of a grid a carver produced (update_num < 1 exactly where sdf == kVal) restores it fully;
 * any other upload is accepted as arbitrary state, but marching cubes then needs update_num, which
 * has no halo planes: on a z slab above the first vcb_extract_iso_begin returns
 * VCB_ERR_UNSUPPORTED. */
int vcb_download_grid(vcb_carver* c, float* sdf, int32_t* update_num);
int vcb_upload_grid(vcb_carver* c, const float* sdf, const int32_t* update_num);
/* Grid checkpoint, no counterpart in the reference (it never serialises the grid): the owned
 * slab's sdf[] and update_num[] with a header (grid size, plane range, carving options) in one raw
 * file.  vcb_load_grid copies the planes that the file and this carver's slab share (any slab
 * partition may restore any other; *planes_loaded tells how many) and refuses files of another grid
 * or other carving options with VCB_ERR_INVALID.  Carving on after a restore gives the same bits as
 * never having stopped.  For a multi-GPU grid vcb_sharded_save_grid writes one file per slab,
 * "<prefix>.z<begin>-<end>.vcbgrid", and vcb_sharded_load_grid restores from a list of files that
 * together cover the grid. */
int vcb_save_grid(vcb_carver* c, const char* path);
int vcb_load_grid(vcb_carver* c, const char* path, int32_t* planes_loaded);
/* DEVICE pointers to the owned slab's state (stream-ordered access only). */
int vcb_grid_device(vcb_carver* c, float** sdf_dev, int32_t** update_num_dev);

/* Measurement aid, no counterpart in the reference: how the fusion kernel treated the (voxel tile,
 * view) pairs of every launch since the last call.  A tile is 32 x 8 x 8 voxels (*voxels_per_tile);
 * pairs[] counts, in this order: general path (sampling with per-voxel range tests), all texels
 * 1.0f (load-free path), all texels kVal (view skipped: every sample would fail the truncation test
 * of voxel_carver.cc:478-480), projection strictly inside the ROI (sampling without range tests),
 * projection outside the ROI (skipped, voxel_carver.cc:462-472).  warp_iterations[] (may be NULL):
 * view iterations the warps really executed on the load-free path [0] and on the sampling paths
 * [1]; one iteration is 256 voxel-views of instructions (32 threads x 8 voxels).  It is less than
 * the pair counts suggest once voxel_max_update_num has retired voxels (voxel_carver.cc:447-450):
 * a warp whose voxels have all reached the cap leaves the view loop; these two are only counted
 * while vcb_set_statistics(c, 1) is in force (off by default: counting costs about 1 % of the
 * fusion kernel's time).  Calling vcb_fusion_class_counts synchronises. */
int vcb_set_statistics(vcb_carver* c, int32_t enabled);
int vcb_fusion_class_counts(vcb_carver* c, uint64_t pairs[5], int64_t* voxels_per_tile,
                            uint64_t warp_iterations[2]);

/* ---- one grid across several GPUs (z slabs) ---------------------------------------------------
 * No counterpart in the reference, which is single-process and single-device: a vcb_sharded stands
 * where ONE vacancy::VoxelCarver stands (voxel_carver.h:95-118) and gives the same results - the
 * slabs together hold exactly the grid, and extract exactly the mesh (vertex numbering and face
 * order of marching_cubes.cc:199-218 over the whole grid), of a single carver.  GPU r owns the z
 * planes [z_bounds[r], z_bounds[r+1]).  Per chunk of <= 64 views, view i is transformed to its 2D
 * SDF (voxel_carver.cc:169-237) by GPU i mod N only; the float images are all-gathered over NVLink
 * and every GPU fuses all of them into its slab (voxel_carver.cc:415-496), the next chunk's
 * transforms and gathers running under the current chunk's fusion.  Marching cubes exchanges the
 * two sdf planes below each slab, the per-slab counts and one plane of vertex ids.
 *
 *   vcb_sharded_create        ONE process drives `n_devices` GPUs (a device may be listed more than
 *                             once: "virtual slabs", used by the single-GPU tests); exchanges are
 *                             peer copies between the devices.
 *   vcb_sharded_create_rank   one process per GPU (`rank` of `world`): exchanges are NCCL calls on
 *                             a communicator built from the 128-byte id that rank 0 obtains from
 *                             vcb_sharded_unique_id and the launcher hands to every rank.  Every
 *                             rank makes the same calls with the same views, vcb_sharded_destroy
 *                             included (only the owner of a view reads its mask).  The ranks map
 *                             each other's image pools (CUDA IPC) and gather with the copy engines;
 *                             NCCL carries the signals, the halo / id planes and the mesh
 *                             (VACANCY_B200_CE_GATHER=0: NCCL all-gathers instead).
 *                             libnccl.so.2 is loaded on first use, never linked.
 *   z_bounds                  NULL (equal slabs) or world + 1 plane indices from 0 to nz; every slab
 *                             holds at least 2 planes.  vcb_sharded_repartition re-creates the slabs
 *                             on other boundaries (grid state is discarded, like Init()). */
#define VCB_UNIQUE_ID_BYTES 128
typedef struct vcb_sharded vcb_sharded;
int vcb_sharded_create(const vcb_options* opt, const int32_t* devices, int32_t n_devices,
                       const int32_t* z_bounds, vcb_sharded** out);
int vcb_sharded_unique_id(void* id128);
int vcb_sharded_create_rank(const vcb_options* opt, int32_t device, int32_t rank, int32_t world,
                            const void* id128, const int32_t* z_bounds, vcb_sharded** out);
void vcb_sharded_destroy(vcb_sharded* s);
const char* vcb_sharded_last_error(const vcb_sharded* s);
int vcb_sharded_info(const vcb_sharded* s, int32_t* world, int32_t* n_local, int32_t dims3[3]);
int vcb_sharded_bounds(const vcb_sharded* s, int32_t* z_bounds /* world + 1 */);
int vcb_sharded_repartition(vcb_sharded* s, const int32_t* z_bounds);
/* Slabs of equal thickness do not cost the same (the ends of the grid see less of the object).  With
 * auto-balance on, the engine places the boundaries itself when the first chunk of views (up to 64)
 * of a fresh grid arrives: it classifies every fusion tile of the whole grid against those views
 * (the classification the fusion kernel uses anyway, milliseconds), sums a cost per layer of 8
 * planes and cuts at equal cost - before anything is fused, so only buffers are re-created.  Call
 * it before the first carve call; handles from vcb_sharded_local are invalid after that first
 * carve call and must be fetched again.  Every rank makes the same call. */
int vcb_sharded_set_auto_balance(vcb_sharded* s, int32_t enabled);
/* The cut itself (host only, no GPU): world + 1 plane indices such that every slab carries about
 * total / world of the layer costs, on whole layers of planes_per_layer planes (the last layer may
 * be short: n_layers = ceil(nz / planes_per_layer)), at least one layer per slab. */
int vcb_balance_bounds(const uint64_t* layer_cost, int32_t n_layers, int32_t planes_per_layer,
                       int32_t nz, int32_t world, int32_t* z_bounds);
/* i-th slab held by THIS process (0 <= i < n_local) and its rank: the carver handle serves
 * vcb_download_grid, vcb_grid_device, vcb_ply_body_*, ... on that slab; do not carve through it. */
int vcb_sharded_local(vcb_sharded* s, int32_t i, vcb_carver** carver, int32_t* rank);
/* The CUDA stream (a cudaStream_t) that orders the i-th local slab's reset, fusion and extraction
 * work - for hosts that time the engine with events or chain their own work behind it. */
int vcb_sharded_stream(vcb_sharded* s, int32_t i, void** main_stream);
/* Init() again (voxel_carver.cc:390-391): every slab back to sdf = kVal, update_num = 0. */
int vcb_sharded_reset(vcb_sharded* s);
/* Carve for n views, as vcb_carve_masks / vcb_carve_masks_device (voxel_carver.cc:394-413).  HOST
 * masks are read during the call only, and only the owner's process / device uploads a mask;
 * DEVICE masks must be readable from the owning GPU (single process: any local device's memory
 * with peer access) and stay valid until the streams have passed this call.  A view with a
 * non-NULL sdf_out entry is fused in a launch of its own so that its 2D SDF can be returned. */
int vcb_sharded_carve_masks(vcb_sharded* s, int32_t n_views, const vcb_view* views,
                            const uint8_t* const* masks, float* const* sdf_out);
int vcb_sharded_carve_masks_device(vcb_sharded* s, int32_t n_views, const vcb_view* views,
                                   const uint8_t* const* masks_dev, float* const* sdf_out_dev);
/* vcb_carve_mask_files for the slabs: every process decodes only the files whose masks it uploads. */
int vcb_sharded_carve_mask_files(vcb_sharded* s, int32_t n_views, const vcb_view* views,
                                 const char* const* paths, int32_t n_threads);
/* Carve(camera, roi, sdf) with ready-made HOST float SDF images (voxel_carver.cc:415-496): every
 * slab fuses them as given. */
int vcb_sharded_carve_sdfs(vcb_sharded* s, int32_t n_views, const vcb_view* views,
                           const float* const* sdfs);
int vcb_sharded_flush(vcb_sharded* s);
int vcb_sharded_synchronize(vcb_sharded* s);
/* ExtractIsoSurface over the whole grid (voxel_carver.cc:540-543): _count returns the GLOBAL sizes
 * on every rank; _fill copies the WHOLE mesh into host buffers - in a single process every device
 * writes its part of them, with one process per GPU the parts are sent to rank 0, whose buffers
 * receive the mesh (the other ranks pass NULL and only send).  _local_mesh: the part of slab i that
 * stays on its device (vcb_extract_iso_device on the slab's carver), ids already global. */
int vcb_sharded_extract_iso_count(vcb_sharded* s, double iso_level, int32_t linear_interp,
                                  int64_t* n_vertices, int64_t* n_faces);
int vcb_sharded_extract_iso_fill(vcb_sharded* s, float* vertices, int32_t* faces);
int vcb_sharded_local_mesh(vcb_sharded* s, int32_t i, int64_t* n_vertices, int64_t* n_faces,
                           int64_t* vertex_offset, int64_t* face_offset);
/* ExtractVoxel over the whole grid (voxel_carver.cc:530-538, extract_voxel.cc:258-317), inside_empty
 * included (the surface test reads the halo plane below each slab): global cube count, then - single
 * process only - all cubes in the reference's raster order into host buffers. */
int vcb_sharded_extract_voxel_count(vcb_sharded* s, int32_t inside_empty, int64_t* n_voxels);
int vcb_sharded_extract_voxel_fill(vcb_sharded* s, float* vertices, int32_t* faces);
int vcb_sharded_save_grid(vcb_sharded* s, const char* prefix);
int vcb_sharded_load_grid(vcb_sharded* s, const char* const* paths, int32_t n_paths);
/* vcb_stage_times of the i-th local slab (its 2D SDF share runs on a side stream, included). */
int vcb_sharded_stage_times(vcb_sharded* s, int32_t i, float ms[4], int64_t launches[4]);

/* Device time in milliseconds spent by this carver's kernels since the last call, by stage:
 * [0] 2D SDF, [1] fusion, [2] ExtractVoxel, [3] marching cubes; launches[4] likewise counts kernel
 * launches. Measured with CUDA events on the carver's stream; calling it synchronises. */
int vcb_stage_times(vcb_carver* c, float ms[4], int64_t launches[4]);

#ifdef __cplusplus
}
#endif
#endif /* VACANCY_B200_H_ */
