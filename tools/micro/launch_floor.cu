// Fixed cost of one GEMM-like launch on B200, built up piece by piece (k3m_gemm_bf16 spends ~7 us
// on a single-tile problem, cuBLAS 3.2 us).  Each variant is launched 20 times back to back in one CUDA graph and
// replayed; the figure is microseconds per launch, dependency latency between graph nodes included.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o launch_floor tools/micro/launch_floor.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

struct Blob { uint8_t b[640]; };           // the size of five CUtensorMap grid constants

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int V>
__global__ void k(const __grid_constant__ Blob blob, uint8_t* g, int* sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  if (V >= 2 && blob.b[threadIdx.x % 640] == 77) *sink = 1;
  if (V >= 3) {
    if (threadIdx.x == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x / 32 == 1) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(512));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  if (V >= 4 && threadIdx.x == 0) {
    const uint8_t* src = g + (size_t)blockIdx.x * 32768;
    for (int i = 0; i < (V >= 5 ? 3 : 1); ++i) {               // V5: three dependent load round trips (12 k-blocks / 5 stages)
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(32768) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(smem_u32(smem)), "l"(src), "r"(32768), "r"(smem_u32(&bar)) : "memory");
      uint32_t ok = 0;
      while (!ok)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(smem_u32(&bar)), "r"(i & 1) : "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 ::"l"(g + (size_t)(256 + blockIdx.x) * 32768), "r"(smem_u32(smem)), "r"(32768) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
  if (V >= 3) {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x / 32 == 1) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(512));
    }
  }
}

template <int V>
void run(const char* what, int grid, int threads, int smem, uint8_t* g, int* sink) {
  Blob blob = {};
  cudaFuncSetAttribute(k<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaStream_t st;
  cudaStreamCreate(&st);
  for (int i = 0; i < 3; ++i) k<V><<<grid, threads, smem, st>>>(blob, g, sink);
  cudaStreamSynchronize(st);
  cudaGraph_t graph;
  cudaGraphExec_t exec;
  cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal);
  for (int i = 0; i < 20; ++i) k<V><<<grid, threads, smem, st>>>(blob, g, sink);
  cudaStreamEndCapture(st, &graph);
  cudaGraphInstantiate(&exec, graph, 0);
  cudaGraphLaunch(exec, st);
  cudaStreamSynchronize(st);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, st);
  for (int i = 0; i < 5; ++i) cudaGraphLaunch(exec, st);
  cudaEventRecord(e1, st);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  printf("V%d grid %3d x %3d thr, %6d B smem: %6.2f us per launch  %-60s (%s)\n", V, grid, threads, smem, ms * 1e3f / 100.f, what,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  uint8_t* g;
  int* sink;
  cudaMalloc(&g, (size_t)512 * 32768);
  cudaMalloc(&sink, 4);
  cudaMemset(g, 0, (size_t)512 * 32768);
  const int big = 224 * 1024;
  for (int grid : {6, 108}) {
    run<0>("empty kernel", grid, 32, 0, g, sink);
    run<1>("empty kernel, GEMM-sized block and shared memory", grid, 320, big, g, sink);
    run<2>("+ 640 B of grid constants (read)", grid, 320, big, g, sink);
    run<3>("+ barrier init, TMEM alloc 512 / dealloc, 2 block syncs", grid, 320, big, g, sink);
    run<4>("+ one 32 KB bulk load, wait, one 32 KB bulk store, wait.read", grid, 320, big, g, sink);
    run<5>("+ three dependent load round trips instead of one", grid, 320, big, g, sink);
    run<5>("same with 576 threads", grid, 576, big, g, sink);
  }
  return 0;
}
