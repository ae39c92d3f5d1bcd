// FFMA vs FFMA2 (fma.rn.f32x2) issue / pipe rate per SM sub-partition on sm_100a: N independent accumulator chains per
// thread, W warps per SMSP (block = 128 * W threads, one block per SM), cycles per warp-instruction from clock64.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_rate tools/micro/ffma2_rate.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int PACKED, int CH>
__global__ void k(float* out, long long* cyc, int iters) {
  float a[CH * 2];
#pragma unroll
  for (int i = 0; i < CH * 2; ++i) a[i] = threadIdx.x * 1e-3f + i;
  const float m = 1.0001f, c = 1e-4f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (PACKED) {
#pragma unroll
      for (int i = 0; i < CH; ++i) {
        unsigned long long r, x, mm, cc;
        asm("mov.b64 %0, {%1, %2};" : "=l"(x) : "f"(a[2 * i]), "f"(a[2 * i + 1]));
        asm("mov.b64 %0, {%1, %2};" : "=l"(mm) : "f"(m), "f"(m));
        asm("mov.b64 %0, {%1, %2};" : "=l"(cc) : "f"(c), "f"(c));
        asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(x), "l"(mm), "l"(cc));
        asm("mov.b64 {%0, %1}, %2;" : "=f"(a[2 * i]), "=f"(a[2 * i + 1]) : "l"(r));
      }
    } else {
#pragma unroll
      for (int i = 0; i < CH * 2; ++i) asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(a[i]) : "f"(a[i]), "f"(m), "f"(c));
    }
  }
  const long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < CH * 2; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int PACKED, int CH>
void run(int warps_per_smsp, float* out, long long* cyc) {
  const int iters = 4096, threads = 128 * warps_per_smsp;
  k<PACKED, CH><<<148, threads>>>(out, cyc, iters);
  k<PACKED, CH><<<148, threads>>>(out, cyc, iters);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double mean = 0;
  for (int i = 0; i < 148; ++i) mean += h[i];
  mean /= 148;
  const double insts = (double)iters * (PACKED ? CH : CH * 2) * warps_per_smsp;     // warp-instructions per SMSP
  printf("%s chains=%d warps/SMSP=%d: %.2f clk per warp-instruction per SMSP, %.1f fp32 FMA lanes/clk/SMSP (%s)\n",
         PACKED ? "FFMA2" : "FFMA ", CH * 2, warps_per_smsp, mean / insts, (double)iters * CH * 2 * warps_per_smsp * 32 / mean,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  float* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4);
  cudaMalloc(&cyc, 148 * 8);
  for (int w : {1, 2, 4, 8}) {
    run<0, 8>(w, out, cyc);
    run<1, 8>(w, out, cyc);
  }
  return 0;
}
