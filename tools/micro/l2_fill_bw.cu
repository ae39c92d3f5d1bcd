// L2 -> shared-memory fill bandwidth of B200, the bound of the GEMM main loops: every CTA
// (one per SM, 1 warp) keeps STAGES bulk copies (cp.async.bulk, CHUNK bytes each, the size of one GEMM k-block)
// in flight from an L2-resident buffer.  Prints aggregate GB/s and bytes per clock and SM for several grid sizes.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2_fill_bw tools/micro/l2_fill_bw.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int STAGES>
__global__ void fill_kernel(const uint8_t* __restrict__ src, size_t src_bytes, int chunk, int iters, int same) {
  extern __shared__ __align__(1024) uint8_t smem_all[];
  __shared__ uint64_t bars_all[4 * STAGES];
  // one issuing thread per warp, each with its own stages and barriers (blockDim.x / 32 issuers per CTA)
  uint8_t* smem = smem_all + (size_t)(threadIdx.x >> 5) * STAGES * chunk;
  uint64_t* bars = bars_all + (threadIdx.x >> 5) * STAGES;
  if ((threadIdx.x & 31) == 0) {
    for (int s = 0; s < STAGES; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const size_t nchunks = src_bytes / chunk;      // a power of two (64 MB buffer): no 64-bit division in the loop
    // same = 1: all CTAs read the same chunk sequence (operand sharing as in a GEMM row of tiles); 0: disjoint
    size_t pos = same ? 0 : (size_t)(blockIdx.x * 4 + (threadIdx.x >> 5)) * 61;
    for (int it = 0; it < iters + STAGES; ++it) {
      const int s = it % STAGES;
      if (it >= STAGES) {                       // wait for the copy that used this stage
        const uint32_t ph = ((it / STAGES) - 1) & 1;
        uint32_t ok = 0;
        while (!ok)
          asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                       : "=r"(ok) : "r"(smem_u32(&bars[s])), "r"(ph) : "memory");
      }
      if (it < iters) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[s])), "r"(chunk) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(smem + (size_t)s * chunk)), "l"(src + (pos & (nchunks - 1)) * chunk), "r"(chunk),
                       "r"(smem_u32(&bars[s])) : "memory");
        pos += 1;
      }
    }
  }
}

template <int STAGES>
void run(const uint8_t* src, size_t bytes, int chunk, int grid, int same, float clk_ghz, int issuers = 1) {
  const int iters = 2000;
  cudaFuncSetAttribute(fill_kernel<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, issuers * STAGES * chunk);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  fill_kernel<STAGES><<<grid, 32 * issuers, issuers * STAGES * chunk>>>(src, bytes, chunk, iters, same);      // warm-up (fills L2)
  cudaEventRecord(e0);
  fill_kernel<STAGES><<<grid, 32 * issuers, issuers * STAGES * chunk>>>(src, bytes, chunk, iters, same);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  const double total = (double)grid * iters * chunk * issuers;
  printf("issuers %d stages %d chunk %5d B grid %3d %s: %8.1f GB/s aggregate, %6.1f GB/s per SM, %5.1f B/clk/SM  (%s)\n", issuers, STAGES, chunk,
         grid, same ? "shared  " : "disjoint", total / ms / 1e6, total / ms / 1e6 / grid, total / ms / 1e6 / grid / clk_ghz,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, 0);
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const float ghz = khz / 1e6f;
  const size_t bytes = 64u << 20;                 // L2-resident (126 MB L2)
  uint8_t* src;
  cudaMalloc(&src, bytes);
  cudaMemset(src, 1, bytes);
  printf("%s, %d SMs, %.3f GHz (max)\n", prop.name, prop.multiProcessorCount, ghz);
  const int grids[] = {1, 8, 37, 74, 111, 148};
  for (int same = 0; same < 2; ++same)
    for (int g : grids) run<5>(src, bytes, 32768, g, same, ghz);
  for (int g : {1, 148}) run<6>(src, bytes, 16384, g, 0, ghz);
  for (int g : {1, 148}) run<3>(src, bytes, 65536, g, 0, ghz);
  for (int g : {1, 148}) run<12>(src, bytes, 16384, g, 0, ghz);
  // several issuing threads per CTA: is the ~375 clocks per copy a per-thread or a per-SM cost?
  for (int g : {1, 148}) run<3>(src, bytes, 16384, g, 0, ghz, 2);
  for (int g : {1, 148}) run<3>(src, bytes, 16384, g, 0, ghz, 4);
  for (int g : {1, 148}) run<3>(src, bytes, 32768, g, 0, ghz, 2);
  for (int g : {1, 148}) run<2>(src, bytes, 8192, g, 0, ghz, 4);
  for (int g : {1, 148}) run<6>(src, bytes, 8192, g, 0, ghz, 1);
  cudaFree(src);
  return 0;
}
