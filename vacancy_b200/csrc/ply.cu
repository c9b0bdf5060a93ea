// F2 - ASCII PLY text of a triangle mesh, produced on the GPU.
//
// Replaces the body of Mesh::WritePly (mesh.cc:583-631) for meshes without vertex colours: one
// line "x y z \n" per vertex (`ofs << float`, i.e. printf("%g") with 6 significant digits) and one
// line "3 i j k \n" per face, byte for byte.  At 1024^3 the mesh has millions of lines and the
// reference's ostream formatting takes seconds; here every line is formatted by one thread with
// exact integer arithmetic (ply_format.cuh), in two passes:
//   ply_len_kernel  : length of every line  -> exclusive scan = byte offset of every line
//   ply_emit_kernel : formats again and writes the characters at the line's offset
// Vertices come first, faces follow (the file layout), so one scan serves both parts.
// Algorithmic bytes: 12 B read + ~25 B written per line; bound by the per-thread formatting work.
#include <stdio.h>
#include <string.h>

#include "common.cuh"
#include "ply_format.cuh"

namespace vcb {

namespace {

__global__ void __launch_bounds__(256) ply_len_kernel(const uint32_t* __restrict__ verts,
                                                      int64_t nv, const int32_t* __restrict__ faces,
                                                      int64_t nf, uint64_t* __restrict__ lens) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nv + nf) return;
  char line[kPlyMaxVertexLine > kPlyMaxFaceLine ? kPlyMaxVertexLine : kPlyMaxFaceLine];
  int n;
  if (i < nv) {
    n = ply_vertex_line(__ldg(verts + 3 * i), __ldg(verts + 3 * i + 1), __ldg(verts + 3 * i + 2),
                        line);
  } else {
    const int64_t j = i - nv;
    n = ply_face_line(__ldg(faces + 3 * j), __ldg(faces + 3 * j + 1), __ldg(faces + 3 * j + 2), line);
  }
  lens[i] = (uint64_t)n;
}

__global__ void __launch_bounds__(256) ply_emit_kernel(const uint32_t* __restrict__ verts,
                                                       int64_t nv, const int32_t* __restrict__ faces,
                                                       int64_t nf,
                                                       const uint64_t* __restrict__ offsets,
                                                       char* __restrict__ text) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nv + nf) return;
  char line[kPlyMaxVertexLine > kPlyMaxFaceLine ? kPlyMaxVertexLine : kPlyMaxFaceLine];
  int n;
  if (i < nv) {
    n = ply_vertex_line(__ldg(verts + 3 * i), __ldg(verts + 3 * i + 1), __ldg(verts + 3 * i + 2),
                        line);
  } else {
    const int64_t j = i - nv;
    n = ply_face_line(__ldg(faces + 3 * j), __ldg(faces + 3 * j + 1), __ldg(faces + 3 * j + 2), line);
  }
  char* __restrict__ out = text + offsets[i];
  for (int k = 0; k < n; k++) out[k] = line[k];
}

}  // namespace

cudaError_t ply_measure(const float* verts_dev, int64_t nv, const int32_t* faces_dev, int64_t nf,
                        PlyText* st, cudaStream_t stream, int64_t* launches) {
  st->nv = nv;
  st->nf = nf;
  st->vertex_bytes = st->face_bytes = 0;
  st->sized = false;
  const int64_t n = nv + nf;
  if (n == 0) {
    st->sized = true;
    return cudaSuccess;
  }
  cudaError_t e = st->lens.reserve((size_t)(n + 2) * sizeof(uint64_t));
  if (e != cudaSuccess) return e;
  uint64_t* lens = static_cast<uint64_t*>(st->lens.ptr);
  ply_len_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(
      reinterpret_cast<const uint32_t*>(verts_dev), nv, faces_dev, nf, lens);
  if (launches) *launches += 1;
  // lens[n] receives the total, lens[0..n) become the offsets
  e = scan_exclusive_u64(lens, lens, n, lens + n, &st->tmp, stream, launches);
  if (e != cudaSuccess) return e;
  uint64_t total = 0, at_faces = 0;
  e = cudaMemcpyAsync(&total, lens + n, sizeof(uint64_t), cudaMemcpyDeviceToHost, stream);
  if (e != cudaSuccess) return e;
  if (nf > 0) {
    e = cudaMemcpyAsync(&at_faces, lens + nv, sizeof(uint64_t), cudaMemcpyDeviceToHost, stream);
    if (e != cudaSuccess) return e;
  }
  e = cudaStreamSynchronize(stream);
  if (e != cudaSuccess) return e;
  st->vertex_bytes = nf > 0 ? at_faces : total;
  st->face_bytes = total - st->vertex_bytes;
  st->sized = true;
  return cudaSuccess;
}

cudaError_t ply_emit(const float* verts_dev, const int32_t* faces_dev, PlyText* st,
                     cudaStream_t stream, int64_t* launches) {
  const int64_t n = st->nv + st->nf;
  const size_t total = (size_t)(st->vertex_bytes + st->face_bytes);
  if (n == 0 || total == 0) return cudaSuccess;
  cudaError_t e = st->text.reserve(total);
  if (e != cudaSuccess) return e;
  ply_emit_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(
      reinterpret_cast<const uint32_t*>(verts_dev), st->nv, faces_dev, st->nf,
      static_cast<const uint64_t*>(st->lens.ptr), static_cast<char*>(st->text.ptr));
  if (launches) *launches += 1;
  return cudaGetLastError();
}

void set_global_error(const char* text);

}  // namespace vcb

// the header of Mesh::WritePly, mesh.cc:596-612 (no vertex colours)
extern "C" int vcb_ply_header(int64_t n_vertices, int64_t n_faces, char* text, size_t capacity,
                              size_t* n_bytes) {
  char buf[256];
  const int n = snprintf(buf, sizeof(buf),
                         "ply\nformat ascii 1.0\nelement vertex %lld\nproperty float x\n"
                         "property float y\nproperty float z\nelement face %lld\n"
                         "property list uchar int vertex_indices\nend_header\n",
                         (long long)n_vertices, (long long)n_faces);
  if (n_bytes) *n_bytes = (size_t)n;
  if (text) {
    if (capacity < (size_t)n) {
      vcb::set_global_error("vcb_ply_header: buffer too small");
      return VCB_ERR_INVALID;
    }
    memcpy(text, buf, (size_t)n);
  }
  return VCB_OK;
}

// Whole file (header + vertex lines + face lines) from HOST arrays: the drop-in for
// Mesh::WritePly's formatting when the mesh is not the carver's own.
extern "C" int vcb_ply_format(int32_t device, const float* vertices, int64_t n_vertices,
                              const int32_t* faces, int64_t n_faces, char* text, size_t capacity,
                              size_t* n_bytes) {
  using namespace vcb;
  char msg[200];
  if (n_vertices < 0 || n_faces < 0 || (n_vertices > 0 && !vertices) || (n_faces > 0 && !faces)) {
    set_global_error("vcb_ply_format: bad argument");
    return VCB_ERR_INVALID;
  }
  size_t header = 0;
  vcb_ply_header(n_vertices, n_faces, nullptr, 0, &header);
  DeviceGuard restore_callers_device;
  cudaError_t e = cudaSetDevice(device);
  float* dv = nullptr;
  int32_t* df = nullptr;
  PlyText st;
  cudaStream_t stream = nullptr;
  int rc = VCB_OK;
  do {
    if (e != cudaSuccess) break;
    if ((e = cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking)) != cudaSuccess) break;
    if (n_vertices > 0) {
      if ((e = cudaMalloc(&dv, (size_t)n_vertices * 12)) != cudaSuccess) break;
      if ((e = cudaMemcpyAsync(dv, vertices, (size_t)n_vertices * 12, cudaMemcpyHostToDevice,
                               stream)) != cudaSuccess)
        break;
    }
    if (n_faces > 0) {
      if ((e = cudaMalloc(&df, (size_t)n_faces * 12)) != cudaSuccess) break;
      if ((e = cudaMemcpyAsync(df, faces, (size_t)n_faces * 12, cudaMemcpyHostToDevice, stream)) !=
          cudaSuccess)
        break;
    }
    if ((e = ply_measure(dv, n_vertices, df, n_faces, &st, stream, nullptr)) != cudaSuccess) break;
    const size_t body = (size_t)(st.vertex_bytes + st.face_bytes);
    if (n_bytes) *n_bytes = header + body;
    if (!text) break;  // size query
    if (capacity < header + body) {
      set_global_error("vcb_ply_format: buffer too small");
      rc = VCB_ERR_INVALID;
      break;
    }
    vcb_ply_header(n_vertices, n_faces, text, capacity, nullptr);
    if ((e = ply_emit(dv, df, &st, stream, nullptr)) != cudaSuccess) break;
    if (body > 0) {
      if ((e = cudaMemcpyAsync(text + header, st.text.ptr, body, cudaMemcpyDeviceToHost, stream)) !=
          cudaSuccess)
        break;
    }
    e = cudaStreamSynchronize(stream);
  } while (0);
  if (stream) {
    cudaStreamSynchronize(stream);
    cudaStreamDestroy(stream);
  }
  cudaFree(dv);
  cudaFree(df);
  st.lens.release();
  st.tmp.release();
  st.text.release();
  if (e != cudaSuccess) {
    snprintf(msg, sizeof(msg), "vcb_ply_format: %s", cudaGetErrorString(e));
    set_global_error(msg);
    return VCB_ERR_CUDA;
  }
  return rc;
}
