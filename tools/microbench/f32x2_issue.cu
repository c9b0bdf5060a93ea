// Microbenchmark (dev tool): issue cost of the sm_100a packed FP32 instructions (FADD2 / FMUL2 /
// FFMA2) against their scalar forms, alone and interleaved with integer work.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -O3 -o f32x2_issue f32x2_issue.cu
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
__device__ __forceinline__ u64 pack(float a, float b) {
  u64 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ float2 unpack(u64 v) {
  float2 r;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
  return r;
}

constexpr int kIters = 4096;
constexpr int kChains = 8;  // independent dependency chains per thread (scalar: 2x as many values)

// mode 0: scalar FADD on 16 values; 1: FADD2 on 8 pairs; 2: scalar FFMA x16; 3: FFMA2 x8;
// 4: FADD2 x8 + 8 IADD; 5: scalar FADD x16 + 8 IADD; 6: FMUL2 x8; 7: 8 IADD only
template <int kMode>
__global__ void __launch_bounds__(256, 4) bench(float* out, float seed, int iseed) {
  float v[2 * kChains];
  u64 p[kChains];
  int q[kChains];
#pragma unroll
  for (int i = 0; i < 2 * kChains; i++) v[i] = seed + threadIdx.x + i;
#pragma unroll
  for (int i = 0; i < kChains; i++) {
    p[i] = pack(v[2 * i], v[2 * i + 1]);
    q[i] = iseed + i + threadIdx.x;
  }
  const float c = seed * 0.5f;
  const u64 cp = pack(c, c);
  for (int it = 0; it < kIters; it++) {
#pragma unroll
    for (int i = 0; i < kChains; i++) {
      if (kMode == 0 || kMode == 5) {
        asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(v[2 * i]) : "f"(c));
        asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(v[2 * i + 1]) : "f"(c));
      }
      if (kMode == 1 || kMode == 4) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(cp));
      if (kMode == 2) {
        asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(v[2 * i]) : "f"(c));
        asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(v[2 * i + 1]) : "f"(c));
      }
      if (kMode == 3) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(p[i]) : "l"(cp));
      if (kMode == 6) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(cp));
      if (kMode == 4 || kMode == 5 || kMode == 7)
        asm volatile("add.s32 %0, %0, %1;" : "+r"(q[i]) : "r"(iseed));
    }
  }
  float acc = 0.0f;
#pragma unroll
  for (int i = 0; i < 2 * kChains; i++) acc += v[i];
#pragma unroll
  for (int i = 0; i < kChains; i++) {
    const float2 u = unpack(p[i]);
    acc += u.x + u.y + (float)q[i];
  }
  if (acc == 12345.678f) out[0] = acc;
}

template <int kMode>
void run(const char* name, float* out, int fp_lane_ops, int instrs) {
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  const int blocks = 148 * 4 * 4;
  bench<kMode><<<blocks, 256>>>(out, 1.0f, 3);
  cudaEventRecord(a);
  bench<kMode><<<blocks, 256>>>(out, 1.0f, 3);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  const double warps = (double)blocks * 8;
  const double winstr = warps * kIters * instrs;
  printf("%-28s %8.3f ms  %7.2f G warp-instr/s  %7.2f T fp32 lane-ops/s\n", name, ms,
         winstr / ms * 1e-6, warps * 32 * kIters * (double)fp_lane_ops / ms * 1e-9);
}

int main() {
  float* out;
  cudaMalloc(&out, 4);
  run<0>("FADD x16", out, 16, 16);
  run<1>("FADD2 x8", out, 16, 8);
  run<2>("FFMA x16", out, 16, 16);
  run<3>("FFMA2 x8", out, 16, 8);
  run<6>("FMUL2 x8", out, 16, 8);
  run<7>("IADD x8", out, 0, 8);
  run<5>("FADD x16 + IADD x8", out, 16, 24);
  run<4>("FADD2 x8 + IADD x8", out, 16, 16);
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
