// bf16 GEMM for sm_100a: TMA -> shared memory (128B swizzle) -> tcgen05.mma (accumulators in
// TMEM) -> tcgen05.ld -> fused epilogue (bias, GELU / GELU', sigmoid, tanh, dropout, residual
// addend, fp32 accumulate for weight gradients with split-K).
//
// One CTA computes one 128 x BN output tile (optionally one K-split of it):
//   warp 0      TMA producer (one elected lane)
//   warp 1      TMEM allocator + MMA issuer (one elected lane)
//   warps 2..5  epilogue: warp w reads TMEM lanes 32*(w%4) .. +31 (one output row per thread)
// Operand layouts ("major" = which dimension is contiguous in global memory):
//   nt  A[M,K]  K-major   B[N,K]  K-major      forward  Y = X W^T
//   nn  A[M,K]  K-major   B[K,N]  MN-major     dgrad    dX = dY W
//   tn  A[K,M]  MN-major  B[K,N]  MN-major     wgrad    dW = dY^T X
// MN-major tiles are loaded as 64(k) x 64(mn) TMA boxes; the UMMA descriptor then uses
// LBO = 8 KB (next 64-wide mn chunk) and SBO = 1 KB (next 8 k-rows).
#include <cuda.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>

#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "common.cuh"

namespace {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 B = one swizzle row

struct GemmParams {
  int M, N, K;
  int kb_per_split;  // k-blocks of one K-split
  int tiles_m, tiles_n, splits;
  void* out;
  int ldc, out_f32, accumulate;
  const bf16* bias;
  int act;
  const bf16* aux_in;
  int ld_aux_in;
  bf16* aux_out;
  int ld_aux_out;
  const void* addend;
  int ld_add, addend_f32;
  DropSpec drop;
  int drop_ld;
  int vec_ok;  // all epilogue pointers 16 B aligned with ld % 8 == 0
  int tma_out; // out (and aux_out) leave through TMA bulk stores
  int fast;    // specialised epilogue (F_* below), 0 = generic
  int probe;     // K3M_GEMM_PROBE, honoured only by a -DK3M_GEMM_PROBE_BUILD library (measurement only, results are
                 // garbage): 1 = no TMA loads, 2 = no MMAs.  The product build carries no probe branch: one extra
                 // predicate per MMA in the issuing thread cost 3 % of GEMM throughput (r01j vs r01i).
  int tin;     // epilogue operand that ARRIVES through TMA into the staging buffer: 0 none, 1 aux_in, 2 addend
  int a3d, b3d;  // MN-major operand described by a 3-D tensor map (see get_map_mn3d): one bulk copy per stage
};

// One problem of a (possibly grouped) launch, in kernel-parameter space: its five tensor maps, its epilogue
// description and where its tiles start in the launch's concatenated tile list.
struct ProblemArgs {
  CUtensorMap map_a, map_b, map_c, map_aux, map_in;
  GemmParams p;
  int tile_begin;
  int total_kb;
};
// MAXP = 1: the single-problem entry point (every index is a compile-time 0).  MAXP = 8: k3m_gemm_grouped_bf16 —
// several independent problems of one layout (the three connection layers of a stage, a text and an image layer,
// a stage's weight gradients) share ONE persistent launch: the tile scheduler walks the concatenated tile list,
// so small problems fill whole waves together instead of each leaving SMs idle behind its own launch latency.
template <int MAXP>
struct GroupArgs {
  int n;
  int total_tiles;
  int snake;      // odd waves walk the tile list backwards (see tile_of in the kernel)
  ProblemArgs prob[MAXP];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Bounded: a protocol bug must surface as a trapped launch (error code at the next API call), never
// as a kernel that spins until the box's watchdog fires.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  uint32_t spins = 0;
  do {
    if (++spins > (1u << 26)) __trap();
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
// ---- CTA-pair (cta_group::2) variants: two CTAs of one cluster (one TPC) work on a 256-row tile; each
// loads its own 128 rows of A and HALF of the B tile, the leader (rank 0) issues one MMA for both.
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// destination in the executing CTA, completion bytes signalled on `bar_cluster` (a shared::cluster address,
// the leader's barrier)
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, uint32_t bar_cluster, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(bar_cluster), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_load_3d_pair(void* dst, const CUtensorMap* map, uint32_t bar_cluster, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
// arrives on the barrier at the same shared-memory offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
// one lane of a converged warp (elect.sync)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}" : "=r"(pred));
  return pred != 0;
}
// No .release.cluster here: what the arrive orders are this warp's TMEM reads, and those are ordered by
// tcgen05.wait::ld + tcgen05.fence::before_thread_sync; a cluster-scope release compiles to MEMBAR.ALL.GPU + ERRBAR
// in front of the arrive — 12-14 % of all stall samples of the K = 768 epilogues (profiles/r03a_gemm_*_hot.txt).
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster) : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t addr, uint32_t r[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, "
      "%15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(addr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_ld16(uint32_t addr, uint32_t r[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(addr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// shared-memory matrix descriptor (SWIZZLE_128B, descriptor version 1)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// ---- epilogue pieces -------------------------------------------------------------------------------
// 8 consecutive bf16 -> 8 floats by integer shifts (bf16 is the high half of an fp32: bit-equal to
// __bfloat1622float2).  The generic epilogue uses this instead of common.cuh's ld8, whose view of a local uint4 through
// a __nv_bfloat162 pointer is type punning: in gemm_kernel<128, 5, 4, ., ., 1, 8> (the grouped 128 x 128 variant, where
// v[] / pre[] live in local memory) the compiler turned the warp-uniform bias load into uniform-register code that read
// registers the load never wrote — elements 3, 5, 7 of every 8-column group came out as garbage
// (tools/debug_group_f32.py, profiles/r03last_group_f32.txt).
__device__ __forceinline__ void ld8_bits(const bf16* p, float v[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  v[0] = __uint_as_float(u.x << 16); v[1] = __uint_as_float(u.x & 0xffff0000u);
  v[2] = __uint_as_float(u.y << 16); v[3] = __uint_as_float(u.y & 0xffff0000u);
  v[4] = __uint_as_float(u.z << 16); v[5] = __uint_as_float(u.z & 0xffff0000u);
  v[6] = __uint_as_float(u.w << 16); v[7] = __uint_as_float(u.w & 0xffff0000u);
}
// values of one 8-column group of one output row: v = final value, pre = GELU pre-activation (aux_out)
__device__ __forceinline__ void compute_group(const GemmParams& p, const uint32_t* acc, int row, int col, int nv,
                                              bool vec, bool has_drop, float v[8], float pre[8],
                                              const float* staged, int tin) {
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(acc[j]);
  if (p.bias) {
    float t[8];
    if (vec) ld8_bits(p.bias + col, t);
    else for (int j = 0; j < 8; ++j) t[j] = j < nv ? ld_f(p.bias + col + j) : 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] += t[j];
  }
  if (p.act == ACT_GELU) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { pre[j] = v[j]; v[j] = gelu_f(v[j]); }
  } else if (p.act == ACT_GELU_GRAD) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { pre[j] = dgelu_f(v[j]); v[j] = gelu_f(v[j]); }
  } else if (p.act == ACT_DGELU || p.act == ACT_DTANH || p.act == ACT_MULAUX) {
    float t[8];
    if (tin == 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j) t[j] = staged[j];
    } else {
      const bf16* ai = p.aux_in + (size_t)row * p.ld_aux_in + col;
      if (vec) ld8_bits(ai, t);
      else for (int j = 0; j < 8; ++j) t[j] = j < nv ? ld_f(ai + j) : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] *= (p.act == ACT_DGELU) ? dgelu_f(t[j]) : (p.act == ACT_MULAUX ? t[j] : (1.0f - t[j] * t[j]));
  } else if (p.act == ACT_SIGMOID) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = sigmoid_f(v[j]);
  } else if (p.act == ACT_TANH) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = tanhf(v[j]);
  }
  if (has_drop) {
    uint32_t w[4];
    const unsigned long long lin = (unsigned long long)row * (unsigned long long)p.drop_ld + (unsigned long long)col;
    rng_words(p.drop.rng, p.drop.site, lin >> 3, w);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] *= drop_mult(p.drop, w, j);
  }
  if (p.addend) {
    float t[8];
    if (tin == 2) {
#pragma unroll
      for (int j = 0; j < 8; ++j) t[j] = staged[j];
    } else if (p.addend_f32) {
      const float* ad = reinterpret_cast<const float*>(p.addend) + (size_t)row * p.ld_add + col;
      if (vec) ld8(ad, t);
      else for (int j = 0; j < 8; ++j) t[j] = j < nv ? ad[j] : 0.f;
    } else {
      const bf16* ad = reinterpret_cast<const bf16*>(p.addend) + (size_t)row * p.ld_add + col;
      if (vec) ld8_bits(ad, t);
      else for (int j = 0; j < 8; ++j) t[j] = j < nv ? ld_f(ad + j) : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] += t[j];
  }
}

// direct (row-per-thread) stores: fallback when the output cannot be described by a TMA tensor map
__device__ __forceinline__ void store_group_direct(const GemmParams& p, int row, int col, int nv, bool vec,
                                                   const float v[8], const float pre[8]) {
  if ((p.act == ACT_GELU || p.act == ACT_GELU_GRAD) && p.aux_out) {
    bf16* ao = p.aux_out + (size_t)row * p.ld_aux_out + col;
    if (vec) st8(ao, pre);
    else for (int j = 0; j < nv; ++j) st_f(ao + j, pre[j]);
  }
  if (p.out_f32) {
    float* o = reinterpret_cast<float*>(p.out) + (size_t)row * p.ldc + col;
    if (p.accumulate) {
      for (int j = 0; j < nv; ++j) atomicAdd(o + j, v[j]);
    } else if (vec) {
      st8(o, v);
    } else {
      for (int j = 0; j < nv; ++j) o[j] = v[j];
    }
  } else {
    bf16* o = reinterpret_cast<bf16*>(p.out) + (size_t)row * p.ldc + col;
    if (vec) st8(o, v);
    else for (int j = 0; j < nv; ++j) st_f(o + j, v[j]);
  }
}

// staging buffer of one epilogue warp: 32 rows x 128 B, SWIZZLE_128B (16-byte chunk c of row r lives at
// chunk c ^ (r & 7)); written row-per-thread without bank conflicts, read by the TMA store engine
// `off` = byte offset inside the (1024-byte aligned) staging buffer in the un-swizzled row-major box
// layout; SWIZZLE_128B XORs the 16-byte chunk index (address bits 4..6) with address bits 7..9
// (SWIZZLE_64B, used for 64-byte rows: chunk bits 4..5 ^= address bits 7..8)
__device__ __forceinline__ uint32_t swz(uint32_t off, uint32_t mask) { return off ^ (((off >> 7) & mask) << 4); }
__device__ __forceinline__ void stage_bf16(uint8_t* buf, uint32_t off, uint32_t mask, const float v[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  *reinterpret_cast<uint4*>(buf + swz(off, mask)) = u;
}
__device__ __forceinline__ void stage_f32(uint8_t* buf, uint32_t off, uint32_t mask, const float v[8]) {
  *reinterpret_cast<float4*>(buf + swz(off, mask)) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(buf + swz(off + 16, mask)) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(map), "r"(smem_u32(src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* map, const void* src, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(map), "r"(smem_u32(src)), "r"(c0), "r"(c1) : "memory");
}

// ---- specialised epilogues ----------------------------------------------------------------------
// The generic epilogue above decides everything per 8-column group at run time (activation, vector
// or scalar access, edge handling): ~35 issued instructions per output element, a third of them
// branches and predicate shuffling — longer than the K = 768 main loop of the tile it overlaps with.
// The four shapes of epilogue the K3M step spends its time in are compiled separately, branch-free:
//   F_PLAIN  out = acc (+ bias) (dropout)                        projections, plain dgrads
//   F_ADD    out = acc (+ bias) (dropout) + addend (via TMA)     attention-out / FFN-down, dgrad + residual
//   F_GELU   aux = acc + bias, out = gelu(aux)                   FFN-up
//   F_DGELU  out = acc * gelu'(aux_in) (aux_in via TMA)          FFN-down dgrad
//   F_ACC    out += acc (fp32, TMA bulk reduce-add)               weight gradients (split-K partial sums)
// Preconditions (host): bf16 output through TMA, no accumulation, N % 8 == 0, 16-byte aligned bias.
enum { F_NONE = 0, F_PLAIN = 1, F_ADD = 2, F_GELU = 3, F_DGELU = 4, F_ACC = 5 };

__device__ __forceinline__ uint4 pack_bf16x8(const float v[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  return u;
}
__device__ __forceinline__ void unpack_bf16x8(const uint4& u, float v[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x; v[2 * i + 1] = f.y;
  }
}

// 16-byte shared-memory accesses by 32-bit shared address (the staging buffers): `*reinterpret_cast<uint4*>(ptr)` on
// a pointer derived from the dynamic shared array compiled to GENERIC ST.E.128 / LD.E.128 with 64-bit address math.
// volatile keeps them ordered against the barrier waits / proxy fences around them (all volatile asm); no "memory"
// clobber, so that ordinary global loads (bias, dropout counter) may still be hoisted across them.
__device__ __forceinline__ void sts16(uint32_t saddr, const uint4& u) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(u.x), "r"(u.y), "r"(u.z), "r"(u.w));
}
__device__ __forceinline__ uint4 lds16(uint32_t saddr) {
  uint4 u;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "r"(saddr));
  return u;
}

// Packed fp32 pairs (Blackwell FFMA2 / FMUL2 / FADD2: two IEEE-rounded fp32 operations per instruction, results equal
// to the scalar fmaf / * / + bit for bit).  The epilogues are bound by instruction issue and dependent-issue latency
// with only four epilogue warps per scheduler; the polynomial part of Phi and the bias / multiply steps halve.
struct F2 { unsigned long long u; };
__device__ __forceinline__ F2 f2(float lo, float hi) {
  F2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r.u) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void f2_get(F2 a, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.u)); }
__device__ __forceinline__ F2 f2_fma(F2 a, F2 b, F2 c) {
  F2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.u) : "l"(a.u), "l"(b.u), "l"(c.u));
  return r;
}
__device__ __forceinline__ F2 f2_mul(F2 a, F2 b) {
  F2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.u) : "l"(a.u), "l"(b.u));
  return r;
}
__device__ __forceinline__ F2 f2_add(F2 a, F2 b) {
  F2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.u) : "l"(a.u), "l"(b.u));
  return r;
}
// gauss_cdf (common.cuh) on a pair: same operations in the same order per element; `e` = exp(-x^2 / 2) of the pair
__device__ __forceinline__ F2 gauss_cdf2(F2 x, F2& e) {
  float x0, x1;
  f2_get(x, x0, x1);
  const F2 z = f2_mul(f2(fabsf(x0), fabsf(x1)), f2(0.70710678118654752f, 0.70710678118654752f));
  float d0, d1;
  f2_get(f2_fma(f2(0.3275911f, 0.3275911f), z, f2(1.0f, 1.0f)), d0, d1);
  const F2 t = f2(rcp_approx(d0), rcp_approx(d1));
  float q0, q1;
  f2_get(f2_mul(f2_mul(x, x), f2(-0.72134752044448170f, -0.72134752044448170f)), q0, q1);
  e = f2(ex2_approx(q0), ex2_approx(q1));
  F2 p = f2_fma(f2(0.5f * 1.061405429f, 0.5f * 1.061405429f), t, f2(0.5f * -1.453152027f, 0.5f * -1.453152027f));
  p = f2_fma(p, t, f2(0.5f * 1.421413741f, 0.5f * 1.421413741f));
  p = f2_fma(p, t, f2(0.5f * -0.284496736f, 0.5f * -0.284496736f));
  p = f2_fma(p, t, f2(0.5f * 0.254829592f, 0.5f * 0.254829592f));
  float h0, h1;
  f2_get(f2_mul(f2_mul(p, t), e), h0, h1);
  return f2(x0 >= 0.0f ? 1.0f - h0 : h0, x1 >= 0.0f ? 1.0f - h1 : h1);
}

// The bias of a 32-column slab (four 16-byte loads) is requested BEFORE the TMEM read of the slab, so that the two
// latencies overlap; loading it per 8-column group right where it is added put a global-load round trip in front of
// every group (long-scoreboard stalls on the unpack instructions, profiles/r03a_gemm_gelu_hot.txt).
__device__ __forceinline__ void load_bias32(const GemmParams& p, int col0, uint4 braw[4]) {
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    const int col = col0 + g * 8;
    braw[g] = (p.bias != nullptr && col < p.N) ? __ldg(reinterpret_cast<const uint4*>(p.bias + col)) : make_uint4(0u, 0u, 0u, 0u);
  }
}
// acc + bias of one 8-column group
__device__ __forceinline__ void acc_bias(const uint32_t* r, const uint4& braw, float v[8]) {
  float t[8];
  unpack_bf16x8(braw, t);
#pragma unroll
  for (int j = 0; j < 8; j += 2)
    f2_get(f2_add(f2(__uint_as_float(r[j]), __uint_as_float(r[j + 1])), f2(t[j], t[j + 1])), v[j], v[j + 1]);
}
// v[0..8) op= t[0..8) on packed pairs
__device__ __forceinline__ void mul8(float v[8], const float t[8]) {
#pragma unroll
  for (int j = 0; j < 8; j += 2) f2_get(f2_mul(f2(v[j], v[j + 1]), f2(t[j], t[j + 1])), v[j], v[j + 1]);
}
__device__ __forceinline__ void add8(float v[8], const float t[8]) {
#pragma unroll
  for (int j = 0; j < 8; j += 2) f2_get(f2_add(f2(v[j], v[j + 1]), f2(t[j], t[j + 1])), v[j], v[j + 1]);
}

// one 32-row x CW-column piece of an output tile, by one epilogue warp (row = lane)
template <int CW, int NBUF, int MODE>
__device__ __forceinline__ void fast_tile(const GemmParams& p, const CUtensorMap* mc, const CUtensorMap* mx,
                                          uint8_t* my_stage, uint32_t& nstore, uint32_t tbase, int row0, int n0,
                                          int lane) {
  constexpr uint32_t RB = CW * 2;                 // bytes per staged row
  constexpr uint32_t SM = RB == 64 ? 3u : 7u;     // 64-byte rows are SWIZZLE_64B boxes
  uint8_t* buf = my_stage + (nstore % NBUF) * 4096;
  uint8_t* abuf = my_stage + ((nstore + 1) % NBUF) * 4096;
  const uint32_t sbuf = smem_u32(buf), sabuf = smem_u32(abuf);
  if (MODE == F_ACC) {
    // Weight-gradient partial sums: the accumulator leaves as fp32 through TMA bulk reduce-add, 16 columns (64-byte
    // rows, SWIZZLE_64B boxes) at a time through the two 2 KB halves of the staging buffer — a half is rewritten
    // only after the reduce issued two chunks earlier has read it (wait_group.read 1), so nothing waits mid-tile.
    // The generic epilogue spent ~35 instructions per element deciding at run time what this path knows statically.
#pragma unroll
    for (int cc = 0; cc < CW; cc += 16) {
      if (n0 + cc >= p.N) break;
      const int h = (cc >> 4) & 1;
      uint32_t r[16];
      tmem_ld16(tbase + (uint32_t)cc, r);
      if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      __syncwarp();
      const uint32_t base = sbuf + (uint32_t)h * 2048u;
#pragma unroll
      for (int q = 0; q < 4; ++q)
        sts16(base + swz((uint32_t)lane * 64u + (uint32_t)(q * 16), 3u), make_uint4(r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]));
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) {
        tma_reduce_add_2d(mc, buf + h * 2048, n0 + cc, row0);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
    nstore += 1;
    return;
  }
  if (MODE == F_PLAIN || MODE == F_GELU) {        // (TMA-fed modes: the operand prefetch waited already)
    if (lane == 0) {
      if (MODE == F_GELU || NBUF == 1) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      else asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    }
    __syncwarp();
  }
  if (MODE == F_GELU && NBUF == 1) {
    // One 4 KB staging buffer, two outputs: the buffer is two 2 KB halves (32 rows x 32 columns, SWIZZLE_64B boxes) —
    // half 0 takes the saved operand of a 32-column slab (pre-activation, or gelu' for ACT_GELU_GRAD), half 1 the
    // activation; each slab reads its accumulator columns ONCE and evaluates Phi once for both outputs.  (The earlier
    // form swept the accumulator twice through the whole buffer: two TMEM reads and two bias adds per element.)
    const bool grad = p.act == ACT_GELU_GRAD;
#pragma unroll
    for (int cc = 0; cc < CW; cc += 32) {
      if (n0 + cc >= p.N) break;
      uint32_t r[32];
      tmem_ld32(tbase + (uint32_t)cc, r);
      if (cc > 0) {                      // the stores of the previous slab must have read both halves
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
      }
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        // (bias per group here, not per slab: this path is the register-hungriest of the kernel and 18 warps cap a
        // thread at 96 registers; the compiler is free to hoist the loads — sts16 carries no memory clobber)
        const int col = n0 + cc + g * 8;
        const uint4 braw = (p.bias != nullptr && col < p.N) ? __ldg(reinterpret_cast<const uint4*>(p.bias + col)) : make_uint4(0u, 0u, 0u, 0u);
        float v[8], a[8];
        acc_bias(r + g * 8, braw, v);
#pragma unroll
        for (int j = 0; j < 8; j += 2) {
          const F2 x = f2(v[j], v[j + 1]);
          F2 e;
          const F2 c = gauss_cdf2(x, e);
          const F2 dg = f2_fma(f2_mul(x, e), f2(0.3989422804014327f, 0.3989422804014327f), c);
          f2_get(grad ? dg : x, a[j], a[j + 1]);
          f2_get(f2_mul(x, c), v[j], v[j + 1]);
        }
        const uint32_t off = swz((uint32_t)lane * 64u + (uint32_t)(g * 16), 3u);
        sts16(sbuf + off, pack_bf16x8(a));
        sts16(sbuf + 2048u + off, pack_bf16x8(v));
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) {
        tma_store_2d(mx, buf, n0 + cc, row0);
        tma_store_2d(mc, buf + 2048, n0 + cc, row0);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      __syncwarp();
    }
    nstore += 1;
    return;
  }
  const bool has_drop = p.drop.thresh != 0;
  const int row = row0 + lane;
  unsigned long long seed = 0ull, step = 0ull;     // read once per tile, not once per 8-column group
  if (MODE != F_GELU && MODE != F_DGELU && has_drop) { seed = p.drop.rng[0]; step = p.drop.rng[1]; }
#pragma unroll
  for (int cc = 0; cc < CW; cc += 32) {
    uint4 braw[4];
    load_bias32(p, n0 + cc, braw);
    uint32_t r[32];
    tmem_ld32(tbase + (uint32_t)cc, r);
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const int col = n0 + cc + g * 8;
      const uint32_t off = swz((uint32_t)lane * RB + (uint32_t)((cc + g * 8) * 2), SM);
      float v[8];
      acc_bias(r + g * 8, braw[g], v);
      if (MODE == F_GELU) {
        if (p.act == ACT_GELU_GRAD) {
          float a[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) { a[j] = dgelu_f(v[j]); v[j] = gelu_f(v[j]); }
          sts16(sabuf + off, pack_bf16x8(a));
        } else {
          sts16(sabuf + off, pack_bf16x8(v));
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = gelu_f(v[j]);
        }
      }
      if (MODE == F_DGELU) {
        float t[8];
        unpack_bf16x8(lds16(sbuf + off), t);
        if (p.act == ACT_MULAUX) {
          mul8(v, t);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] *= dgelu_f(t[j]);
        }
      }
      if (MODE != F_GELU && MODE != F_DGELU && has_drop) {
        uint32_t w[4];
        const unsigned long long lin = (unsigned long long)row * (unsigned long long)p.drop_ld + (unsigned long long)col;
        const unsigned long long idx = lin >> 3;      // == rng_words(p.drop.rng, p.drop.site, idx, w)
        philox4x32((uint32_t)idx, (uint32_t)(idx >> 32), p.drop.site, (uint32_t)step, (uint32_t)seed,
                   (uint32_t)(seed >> 32), w);
        float dm[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) dm[j] = drop_mult(p.drop, w, j);
        mul8(v, dm);
      }
      if (MODE == F_ADD) {
        float t[8];
        unpack_bf16x8(lds16(sbuf + off), t);
        add8(v, t);
      }
      sts16(sbuf + off, pack_bf16x8(v));
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();
  if (lane == 0) {
    tma_store_2d(mc, buf, n0, row0);
    if (MODE == F_GELU) tma_store_2d(mx, abuf, n0, row0);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  }
  nstore += MODE == F_GELU ? 2 : 1;
}

// Persistent, warp-specialised: every CTA walks tiles t = blockIdx.x, blockIdx.x + gridDim.x, ...
// of the launch's tile list (problem after problem; inside a problem tile = (split, m, n), n fastest so
// neighbouring CTAs share the A panel in L2).  The smem ring (TMA -> MMA) runs across tile boundaries and the
// accumulator is double-buffered in TMEM, so the epilogue of tile i overlaps the main loop of tile i + 1 —
// also when the two tiles belong to different problems of a grouped launch.
// STAGES == 2 marks the "dual" variant: a 2-stage ring and one staging buffer per epilogue warp keep a CTA
// under 100 KB of shared memory and 256 TMEM columns, so TWO of them (of the same or of different launches)
// share an SM — for the single-wave M = 2304 GEMMs that run concurrently on several streams.
template <int BN, int STAGES, int EW, bool A_MN, bool B_MN, int CTAS, int MAXP>
__global__ void __launch_bounds__(64 + 128 * EW, STAGES <= 2 ? 2 : 1)
gemm_kernel(const __grid_constant__ GroupArgs<MAXP> g) {
  // warps: 0 = TMA producer, 1 = MMA issuer, 2 .. 2+4*EW-1 = epilogue.  Epilogue warp e handles the
  // TMEM lane quarter (warp % 4) and the column group e / 4 (CW = BN / EW columns) of every tile.
  // CTAS == 2: CTA pair, 256 x BN tile per pair, this CTA holds rows rank*128.. of A / D and columns
  // rank*BN/2.. of B (see the cta_group::2 helpers above)
  constexpr int BNL = BN / CTAS;              // B columns this CTA loads
  constexpr int A_BYTES = BM * BK * 2;
  constexpr int B_BYTES = BNL * BK * 2;
  constexpr int CW = BN / EW;
  static_assert(CW == 32 || CW == 64, "an epilogue warp owns 32 or 64 output columns");
  constexpr int NBUF = (EW <= 2 && STAGES > 2) ? 2 : 1;        // staging buffers (4 KB each) per epilogue warp
  constexpr uint32_t TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_BYTES;
  uint8_t* sStage = sB + STAGES * B_BYTES;                                   // [4*EW warps][NBUF][4096], 1024-aligned
  uint64_t* full = reinterpret_cast<uint64_t*>(sStage + 4 * EW * NBUF * 4096);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;   // [2]
  uint64_t* in_bars = tmem_empty + 2;     // [4 * EW] one per epilogue warp: TMA-fed epilogue operand landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(in_bars + 4 * EW);

  // the broadcast tells the compiler that `warp` is warp-uniform: the role branches below are then uniform control
  // flow and the MMA warp's addresses / descriptors live in uniform registers (see the MMA branch)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int total_tiles = g.total_tiles;
  const uint32_t rank = CTAS == 2 ? cluster_ctarank() : 0u;
  const bool leader = rank == 0;
  const int tile0 = blockIdx.x / CTAS, tstride = gridDim.x / CTAS;     // tiles are walked per CTA pair
  // Wave w of the walk: tile w * tstride + tile0, and with `snake` the odd waves mirrored (tstride - 1 - tile0).  The
  // tile list is ordered longest-k first, so the last waves of a grouped launch mix long and short tiles: the plain
  // stride hands the same low-numbered CTAs the long tile of EVERY mixed wave (e.g. 108 tiles of 48 k-blocks + 76
  // of 16 on 74 pairs: 48+48+16 against an average of 86), the mirrored walk alternates (48+16+16 / 48+48).
  const int tile0m = g.snake ? tstride - 1 - tile0 : tile0;
  auto tile_of = [&](int w) { return w * tstride + ((w & 1) ? tile0m : tile0); };
  // problem of a tile of the launch's list: tiles only grow along a CTA's walk, so `pi` only moves forward
  auto locate = [&](int tile, int& pi) {
    if (MAXP > 1) {
      while (pi + 1 < g.n && tile >= g.prob[pi + 1].tile_begin) ++pi;
    }
  };

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 4 * EW * CTAS); // one arrive per epilogue warp (of both CTAs of a pair)
    }
    for (int s = 0; s < 4 * EW; ++s) mbar_init(&in_bars[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    if (CTAS == 2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                   "r"(TMEM_COLS));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                   "r"(TMEM_COLS));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  if (CTAS == 2) cluster_sync_all();     // the peer's barriers must be initialised before anything signals them
  else __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  // Programmatic dependent launch: everything above (barrier init, TMEM allocation, cluster sync) touches no
  // global memory and may run while the previous kernel of the stream drains; from here on its results are
  // needed.  Both instructions are no-ops when the launch did not carry the attribute.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

  if (warp == 0) {
    if (lane == 0) {
      for (int i = 0; i < (MAXP > 1 ? g.n : 1); ++i) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&g.prob[i].map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&g.prob[i].map_b) : "memory");
      }
      uint32_t it = 0;
      int pi = 0;
      const uint32_t full0 = CTAS == 2 ? mapa_u32(smem_u32(&full[0]), 0) : 0u;   // the leader's full barriers
      for (int w = 0, tile; (tile = tile_of(w)) < total_tiles; ++w) {
        locate(tile, pi);
        const ProblemArgs& pr = g.prob[MAXP > 1 ? pi : 0];
        const GemmParams& p = pr.p;
        const CUtensorMap* map_a = &pr.map_a;
        const CUtensorMap* map_b = &pr.map_b;
        const int lt_ = tile - pr.tile_begin;
        const int tiles_mn = p.tiles_m * p.tiles_n;
        const int split = lt_ / tiles_mn, mn = lt_ % tiles_mn;
        const int m0 = (mn / p.tiles_n) * (BM * CTAS) + (int)rank * BM, n0 = (mn % p.tiles_n) * BN + (int)rank * BNL;
        const int kb0 = split * p.kb_per_split;
        int nkb = pr.total_kb - kb0;
        if (nkb > p.kb_per_split) nkb = p.kb_per_split;
        for (int i = 0; i < nkb; ++i, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(&empty[s], ph ^ 1);
#ifdef K3M_GEMM_PROBE_BUILD
          if (p.probe == 1) {                    // bottleneck bisect: stages turn "full" without any bytes moving
            if (leader) mbar_expect_tx(&full[s], 0);
            continue;
          }
#endif
          if (leader) mbar_expect_tx(&full[s], CTAS * (A_BYTES + B_BYTES));
          const int k0 = (kb0 + i) * BK;
          uint8_t* a = sA + s * A_BYTES;
          uint8_t* b = sB + s * B_BYTES;
          if (CTAS == 2) {
            const uint32_t bar = full0 + (uint32_t)s * 8u;
            if (A_MN) {
              if (p.a3d) {
                tma_load_3d_pair(a, map_a, bar, 0, k0, m0 >> 6);
              } else {
#pragma unroll
                for (int c = 0; c < BM / 64; ++c) tma_load_2d_pair(a + c * 8192, map_a, bar, m0 + c * 64, k0);
              }
            } else {
              tma_load_2d_pair(a, map_a, bar, k0, m0);
            }
            if (B_MN) {
              if (p.b3d) {
                tma_load_3d_pair(b, map_b, bar, 0, k0, n0 >> 6);
              } else {
#pragma unroll
                for (int c = 0; c < BNL / 64; ++c) tma_load_2d_pair(b + c * 8192, map_b, bar, n0 + c * 64, k0);
              }
            } else {
              tma_load_2d_pair(b, map_b, bar, k0, n0);
            }
          } else {
            if (A_MN) {
              if (p.a3d) {
                tma_load_3d(a, map_a, &full[s], 0, k0, m0 >> 6);
              } else {
#pragma unroll
                for (int c = 0; c < BM / 64; ++c) tma_load_2d(a + c * 8192, map_a, &full[s], m0 + c * 64, k0);
              }
            } else {
              tma_load_2d(a, map_a, &full[s], k0, m0);
            }
            if (B_MN) {
              if (p.b3d) {
                tma_load_3d(b, map_b, &full[s], 0, k0, n0 >> 6);
              } else {
#pragma unroll
                for (int c = 0; c < BN / 64; ++c) tma_load_2d(b + c * 8192, map_b, &full[s], n0 + c * 64, k0);
              }
            } else {
              tma_load_2d(b, map_b, &full[s], k0, n0);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // All 32 lanes walk the loops and wait on the barriers; one elected lane issues.  Issuing from inside an
    // `if (lane == 0)` region made every operand of tcgen05.mma a per-thread value: the compiler moved five of them
    // into uniform registers with an ELECT / R2UR / branch loop before EACH MMA (~20 instructions, ~150 clocks —
    // longer than a 128 x 128 x 16 MMA runs).
    if (leader) {
      const bool issuer = elect_one();
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((A_MN ? 1u : 0u) << 15) |
                                 ((B_MN ? 1u : 0u) << 16) | ((uint32_t)(BN >> 3) << 17) |
                                 ((uint32_t)((BM * CTAS) >> 4) << 24);
      uint32_t it = 0, lt = 0;
      int pi = 0;
      for (int tile; (tile = tile_of((int)lt)) < total_tiles; ++lt) {
        locate(tile, pi);
        const ProblemArgs& pr = g.prob[MAXP > 1 ? pi : 0];
        const int tiles_mn = pr.p.tiles_m * pr.p.tiles_n;
        const int split = (tile - pr.tile_begin) / tiles_mn;
        const int kb0 = split * pr.p.kb_per_split;
        int nkb = pr.total_kb - kb0;
        if (nkb > pr.p.kb_per_split) nkb = pr.p.kb_per_split;
        const uint32_t as = lt & 1, aph = (lt >> 1) & 1;
        mbar_wait(&tmem_empty[as], aph ^ 1);          // epilogue (of both CTAs) has drained this accumulator
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t tacc = tmem_base + as * BN;
        for (int i = 0; i < nkb; ++i, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(&full[s], ph);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t a = smem_u32(sA + s * A_BYTES), b = smem_u32(sB + s * B_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = A_MN ? make_desc(a + k * 2048, 8192, 1024) : make_desc(a + k * 32, 0, 1024);
            const uint64_t db = B_MN ? make_desc(b + k * 2048, 8192, 1024) : make_desc(b + k * 32, 0, 1024);
#ifdef K3M_GEMM_PROBE_BUILD
            if (pr.p.probe == 2) continue;       // bottleneck bisect: the ring drains at the pace of the loads alone
#endif
            if (issuer) {
              if (CTAS == 2) umma_bf16_pair(tacc, da, db, idesc, (i > 0 || k > 0) ? 1u : 0u);
              else umma_bf16(tacc, da, db, idesc, (i > 0 || k > 0) ? 1u : 0u);
            }
          }
          if (issuer) {
            if (CTAS == 2) umma_commit_pair(&empty[s]);
            else umma_commit(&empty[s]);
          }
          __syncwarp();
        }
        if (issuer) {
          if (CTAS == 2) umma_commit_pair(&tmem_full[as]);
          else umma_commit(&tmem_full[as]);
        }
        __syncwarp();
      }
    }
  } else {
    // ---- epilogue warps: one output row per thread, CW columns per warp ----
    // Outputs leave through a per-warp swizzled staging buffer and TMA (bulk store, or bulk reduce-add
    // for split-K / accumulated weight gradients): full-line coalesced writes whatever the row stride.
    // A bf16 epilogue operand (GELU' pre-activation or the residual addend) ARRIVES the same way: its
    // 32 x CW box is bulk-loaded into the staging buffer while the main loop of the tile is still
    // running, read row-per-thread from shared memory and overwritten in place by the result
    // (row-per-thread global loads cost 32 L1 wavefronts per instruction and made these epilogues
    // longer than the K = 768 main loop).
    const int e = warp - 2, q = warp & 3, cg = e >> 2;
    uint8_t* my_stage = sStage + e * NBUF * 4096;
    uint64_t* in_bar = in_bars + e;
    uint32_t in_phase = 0;
    uint32_t nstore = 0;
    if (lane == 0) {
      for (int i = 0; i < (MAXP > 1 ? g.n : 1); ++i) {
        const ProblemArgs& pr = g.prob[i];
        if (!pr.p.tma_out) continue;
        asm volatile("prefetch.tensormap [%0];" ::"l"(&pr.map_c) : "memory");
        if ((pr.p.act == ACT_GELU || pr.p.act == ACT_GELU_GRAD) && pr.p.aux_out != nullptr) asm volatile("prefetch.tensormap [%0];" ::"l"(&pr.map_aux) : "memory");
        if (pr.p.tin) asm volatile("prefetch.tensormap [%0];" ::"l"(&pr.map_in) : "memory");
      }
    }
    // bulk-load the epilogue operand box of `tile` (of problem `pj`) into the buffer its result will be staged in
    auto prefetch_in = [&](int tile, int pj) {
      const ProblemArgs& pr = g.prob[MAXP > 1 ? pj : 0];
      if (!pr.p.tin) return;
      const int tiles_mn = pr.p.tiles_m * pr.p.tiles_n;
      const int mn = (tile - pr.tile_begin) % tiles_mn;
      const int m0 = (mn / pr.p.tiles_n) * (BM * CTAS) + (int)rank * BM, n0 = (mn % pr.p.tiles_n) * BN + cg * CW;
      if (n0 >= pr.p.N) return;
      if (lane == 0) {
        // the bulk store that last read this buffer must have finished reading it
        if (NBUF == 1) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        else asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        mbar_expect_tx(in_bar, 32 * CW * 2);
        tma_load_2d(my_stage + (nstore % NBUF) * 4096, &pr.map_in, in_bar, n0, m0 + q * 32);
      }
    };
    int pi = 0, pnext = 0;
    if (tile_of(0) < total_tiles) {
      locate(tile_of(0), pnext);
      prefetch_in(tile_of(0), pnext);
    }
    const uint32_t tmem_empty0 = CTAS == 2 ? mapa_u32(smem_u32(&tmem_empty[0]), 0) : 0u;   // the leader's
    uint32_t lt = 0;
    for (int tile; (tile = tile_of((int)lt)) < total_tiles; ++lt) {
      locate(tile, pi);
      const ProblemArgs& pr = g.prob[MAXP > 1 ? pi : 0];
      const GemmParams& p = pr.p;
      const CUtensorMap* map_c = &pr.map_c;
      const CUtensorMap* map_aux = &pr.map_aux;
      const bool has_drop = p.drop.thresh != 0;
      const bool want_aux = (p.act == ACT_GELU || p.act == ACT_GELU_GRAD) && p.aux_out != nullptr;
      const int esz = p.out_f32 ? 4 : 2;
      const int chunk_cols = (CW * esz > 128) ? 128 / esz : CW;      // columns per staged box (<= 128 B per row)
      const uint32_t row_bytes = (uint32_t)(chunk_cols * esz);
      const uint32_t smask = row_bytes == 64 ? 3u : 7u, amask = CW * 2 == 64 ? 3u : 7u;   // 64 B rows use SWIZZLE_64B
      const int tin = p.tin;                                          // host: only with tma_out, bf16 out, !want_aux
      const int tiles_mn = p.tiles_m * p.tiles_n;
      const int mn = (tile - pr.tile_begin) % tiles_mn;
      const int m0 = (mn / p.tiles_n) * (BM * CTAS) + (int)rank * BM, n0 = (mn % p.tiles_n) * BN + cg * CW;
      const uint32_t as = lt & 1, aph = (lt >> 1) & 1;
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < p.M;
      mbar_wait(&tmem_full[as], aph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN + cg * CW;
      if (n0 < p.N) {
        if (tin) {
          mbar_wait(in_bar, in_phase);
          in_phase ^= 1;
        }
        uint8_t* buf = nullptr;
        uint8_t* abuf = nullptr;
        if (p.fast) {
          const int row0 = m0 + q * 32;
          switch (p.fast) {
            case F_PLAIN: fast_tile<CW, NBUF, F_PLAIN>(p, map_c, map_aux, my_stage, nstore, tbase, row0, n0, lane); break;
            case F_ADD: fast_tile<CW, NBUF, F_ADD>(p, map_c, map_aux, my_stage, nstore, tbase, row0, n0, lane); break;
            case F_GELU: fast_tile<CW, NBUF, F_GELU>(p, map_c, map_aux, my_stage, nstore, tbase, row0, n0, lane); break;
            case F_ACC: fast_tile<CW, NBUF, F_ACC>(p, map_c, map_aux, my_stage, nstore, tbase, row0, n0, lane); break;
            default: fast_tile<CW, NBUF, F_DGELU>(p, map_c, map_aux, my_stage, nstore, tbase, row0, n0, lane); break;
          }
        } else {
#pragma unroll 1
          for (int cc = 0; cc < CW; cc += 32) {
            if (!p.tma_out && n0 + cc >= p.N) break;
            const int lc = cc % chunk_cols;            // column of this 32-wide slab inside the staged box
            if (p.tma_out && lc == 0) {
              buf = my_stage + (nstore % NBUF) * 4096;
              abuf = my_stage + ((nstore + 1) % NBUF) * 4096;
              // the bulk store that last read the buffer(s) must have finished reading (with a TMA-fed
              // operand the prefetch already waited for that before it overwrote the buffer)
              if (lane == 0 && !tin) {
                if (want_aux || NBUF == 1) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                else asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
              }
              __syncwarp();
            }
            uint32_t r[32];
            tmem_ld32(tbase + (uint32_t)cc, r);
#pragma unroll
            for (int gq = 0; gq < 4; ++gq) {
              const int col = n0 + cc + gq * 8;
              const int nv = (p.N - col) < 8 ? (p.N - col) : 8;
              const bool ok = row_ok && nv > 0;
              const bool vec = p.vec_ok && nv == 8;
              const uint32_t off = (uint32_t)lane * row_bytes + (uint32_t)((lc + gq * 8) * esz);
              float v[8], pre[8], staged[8];
              if (tin) {
                ld8_bits(reinterpret_cast<const bf16*>(buf + swz(off, smask)), staged);
              }
              if (ok) {
                compute_group(p, r + gq * 8, row, col, nv, vec, has_drop, v, pre, staged, tin);
              } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = pre[j] = 0.f;
              }
              if (p.tma_out) {
                {
                  if (p.out_f32) stage_f32(buf, off, smask, v);
                  else stage_bf16(buf, off, smask, v);
                  if (want_aux) {
                    if (NBUF > 1) {
                      stage_bf16(abuf, (uint32_t)lane * (uint32_t)(CW * 2) + (uint32_t)((cc + gq * 8) * 2), amask, pre);
                    } else if (ok) {   // fp32 out + pre-activation with one buffer: row-per-thread store
                      bf16* ao = p.aux_out + (size_t)row * p.ld_aux_out + col;
                      if (vec) st8(ao, pre);
                      else for (int j = 0; j < nv; ++j) st_f(ao + j, pre[j]);
                    }
                  }
                }
              } else if (ok) {
                store_group_direct(p, row, col, nv, vec, v, pre);
              }
            }
            if (p.tma_out && (lc + 32 == chunk_cols)) {
              asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
              __syncwarp();
              if (lane == 0) {
                const int c0 = n0 + cc + 32 - chunk_cols;
                if (p.accumulate) tma_reduce_add_2d(map_c, buf, c0, m0 + q * 32);
                else tma_store_2d(map_c, buf, c0, m0 + q * 32);
                if (want_aux && NBUF > 1) tma_store_2d(map_aux, abuf, n0, m0 + q * 32);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
              }
              nstore += (want_aux && NBUF > 1) ? 2 : 1;
            }
          }
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) {
        if (CTAS == 2) mbar_arrive_cluster(tmem_empty0 + as * 8u);
        else asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&tmem_empty[as])) : "memory");
      }
      const int tnext = tile_of((int)lt + 1);
      if (tnext < total_tiles) {
        locate(tnext, pnext);
        prefetch_in(tnext, pnext);
      }
    }
    // the staging buffers must outlive the bulk stores' READS only; global visibility of the writes is
    // given by grid completion
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  if (CTAS == 2) cluster_sync_all();     // the leader's MMAs read the peer's shared memory and write its TMEM
  else __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (CTAS == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS));
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS));
  }
}


// ---- host side ------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(f);
  });
  return fn;
}

typedef std::tuple<const void*, int, int, int, int, int, int> MapKey;
std::map<MapKey, CUtensorMap> g_maps;
std::mutex g_maps_mu;

// 2-D bf16 tensor [outer, inner] with row stride ld (elements), box [box_outer, box_inner], SW128
bool get_map(const void* ptr, int inner, int outer, int ld, int box_inner, int box_outer, CUtensorMap* out,
             int esize = 2) {
  MapKey key(ptr, inner, outer, ld, box_inner, box_outer, esize);
  {
    std::lock_guard<std::mutex> lk(g_maps_mu);
    auto it = g_maps.find(key);
    if (it != g_maps.end()) {
      *out = it->second;
      return true;
    }
  }
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    k3m_set_error("cuTensorMapEncodeTiled entry point not available");
    return false;
  }
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)ld * (cuuint64_t)esize};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMap m;
  CUresult r = enc(&m, esize == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   box_inner * esize == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    k3m_set_error("cuTensorMapEncodeTiled failed (%d) ptr=%p inner=%d outer=%d ld=%d box=%dx%d", (int)r, ptr, inner,
                  outer, ld, box_inner, box_outer);
    return false;
  }
  std::lock_guard<std::mutex> lk(g_maps_mu);
  if (g_maps.size() > 65536) g_maps.clear();
  g_maps[key] = m;
  *out = m;
  return true;
}

// MN-major operand [K, MN] (row stride ld) as a 3-D tensor (mn % 64, k, mn / 64) with strides (1, ld, 64): a box
// (64, 64, chunks) lands in shared memory exactly like `chunks` separate 64 x 64 boxes 8 KB apart — the layout the
// MN-major UMMA descriptors expect — but is ONE bulk copy.  The issuing thread manages one bulk copy per ~250 clocks
// whatever its size (profiles/r01i_l2_fill_bw.txt): the four 8 KB copies per stage of the tn layout took ~1000 clocks
// against ~820 for the stage's MMAs, which is why weight gradients topped out near 1000 TF/s.  Needs MN % 64 == 0
// (the columns of the last chunk would otherwise alias the next row); failure falls back to the 2-D boxes.
typedef std::tuple<const void*, int, int, int, int> Map3Key;
std::map<Map3Key, CUtensorMap> g_maps3;
bool get_map_mn3d(const void* ptr, int mn, int k, int ld, int chunks, CUtensorMap* out) {
  static const int off = getenv("K3M_GEMM_NO_3D") ? atoi(getenv("K3M_GEMM_NO_3D")) : 0;
  if (off || mn % 64 != 0) return false;
  Map3Key key(ptr, mn, k, ld, chunks);
  {
    std::lock_guard<std::mutex> lk(g_maps_mu);
    auto it = g_maps3.find(key);
    if (it != g_maps3.end()) {
      *out = it->second;
      return true;
    }
  }
  EncodeTiledFn enc = get_encode();
  if (!enc) return false;
  cuuint64_t dims[3] = {64, (cuuint64_t)k, (cuuint64_t)(mn / 64)};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, 128};
  cuuint32_t box[3] = {64, 64, (cuuint32_t)chunks};
  cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return false;
  std::lock_guard<std::mutex> lk(g_maps_mu);
  if (g_maps3.size() > 65536) g_maps3.clear();
  g_maps3[key] = m;
  *out = m;
  return true;
}

int g_sm_limit = 0;      // k3m_gemm_set_sm_limit
int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  }
  return (g_sm_limit > 1 && g_sm_limit < n) ? (g_sm_limit & ~1) : n;
}

template <int BN, int STAGES, int EW, bool A_MN, bool B_MN, int CTAS, int MAXP>
int launch(GroupArgs<MAXP>& g, cudaStream_t st) {
  // operand ring + epilogue staging (4*EW warps x NBUF x 4 KB) + barriers + 1 KB alignment slack
  constexpr int NBUF = (EW <= 2 && STAGES > 2) ? 2 : 1;
  constexpr int SMEM = STAGES * (BM * BK * 2 + (BN / CTAS) * BK * 2) + 4 * EW * NBUF * 4096 +
                       (2 * STAGES + 4 + 4 * EW) * 8 + 16 + 1024;
  static_assert(SMEM <= 227 * 1024, "shared memory budget");
  static bool configured = false;
  auto kern = gemm_kernel<BN, STAGES, EW, A_MN, B_MN, CTAS, MAXP>;
  if (!configured) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM) != cudaSuccess) {
      k3m_set_error("gemm: cannot set dynamic shared memory to %d", SMEM);
      return K3M_ERR_CUDA;
    }
    configured = true;
  }
  int total = 0;
  for (int i = 0; i < g.n; ++i) {
    GemmParams& p = g.prob[i].p;
    p.tiles_m = (p.M + BM * CTAS - 1) / (BM * CTAS);
    p.tiles_n = (p.N + BN - 1) / BN;
    g.prob[i].tile_begin = total;
    total += p.tiles_m * p.tiles_n * p.splits;
  }
  g.total_tiles = total;
  static const int snake = getenv("K3M_GEMM_SNAKE") ? atoi(getenv("K3M_GEMM_SNAKE")) : 1;
  g.snake = snake;
  const int slots = num_sms() * (STAGES <= 2 ? 2 : 1) / CTAS;     // CTAs (or CTA pairs) resident at once
  const int grid = (total < slots ? total : slots) * CTAS;
  static const int pdl = getenv("K3M_GEMM_PDL") ? atoi(getenv("K3M_GEMM_PDL")) : 0;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(64 + 128 * EW);
  cfg.dynamicSmemBytes = SMEM;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (CTAS == 2) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = CTAS;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  if (cudaLaunchKernelEx(&cfg, kern, g) != cudaSuccess) {
    k3m_set_error("gemm: launch failed: %s", cudaGetErrorString(cudaGetLastError()));
    return K3M_ERR_CUDA;
  }
  return k3m_check_launch("k3m_gemm_bf16");
}

// (tile width, epilogue column groups) variants: 256/4 (16 epilogue warps), 128/4, 128/2, 64/1; `ctas` = 2
// selects the CTA-pair kernel (256 x 256 tile per pair, 5-stage ring of 32 KB stages)
template <bool A_MN, bool B_MN>
int dispatch(int bn, int ew, int ctas, GroupArgs<1>& g, cudaStream_t st) {
  if (bn == 64) return launch<64, 6, 1, A_MN, B_MN, 1, 1>(g, st);
  if (bn == 256 && ctas == 2) return launch<256, 5, 4, A_MN, B_MN, 2, 1>(g, st);
  if (bn == 128 && ctas == 2) return launch<128, 6, 4, A_MN, B_MN, 2, 1>(g, st);
  if (bn == 256) return launch<256, 3, 4, A_MN, B_MN, 1, 1>(g, st);
  if (ew == 0) return launch<128, 2, 2, A_MN, B_MN, 1, 1>(g, st);      // dual: two CTAs per SM
  if (ew == 4) return launch<128, 5, 4, A_MN, B_MN, 1, 1>(g, st);
  return launch<128, 5, 2, A_MN, B_MN, 1, 1>(g, st);
}
// grouped launches: the CTA-pair kernel, or 128 x 128 single-CTA tiles for problems too small for it
template <bool A_MN, bool B_MN>
int dispatch_group(int ctas, GroupArgs<8>& g, cudaStream_t st) {
  if (ctas == 2) return launch<256, 5, 4, A_MN, B_MN, 2, 8>(g, st);
  return launch<128, 5, 4, A_MN, B_MN, 1, 8>(g, st);
}

// tile width: 128 x 256 (85 flop per operand byte from L2) once it still gives every SM a tile, else
// 128 x 128; operand traffic from L2, not the tensor pipe, bounds these K <= 3072 problems
int pick_bn(int M, int N, int K, int force) {
  if (N <= 64) return 64;
  if (force == 64 || force == 128 || force == 256) return force;
  if (K >= 4096 && N >= 256) return 256;
  const long tiles256 = (long)((M + BM - 1) / BM) * ((N + 255) / 256);
  return (N >= 256 && tiles256 >= (long)num_sms() * 3 / 4) ? 256 : 128;
}

// Split-K plan of the accumulating (weight-gradient) problems of one launch.  The persistent walk hands tile t of the
// list (problems ordered longest tile first, split-major inside a problem) to pair t % P of wave t / P, mirrored on odd
// waves; a tile costs its k-blocks plus a fixed part (accumulator hand-over, fp32 reduce-add of a 128 KB tile).
// Every "target k-blocks per tile" L gives one candidate plan (splits_i = round(kb_i / L), at least 4 k-blocks per
// split); the plan with the smallest modelled makespan wins.  The earlier rule (double the splits until the launch
// has three waves of tiles) ignored how the tile lengths of a mixed group fall onto the waves.
double model_makespan(int n, const long* base_tiles, const int* kb, const int* splits, int P, bool snake) {
  int ord[8];
  int kbs[8];
  for (int i = 0; i < n; ++i) { ord[i] = i; kbs[i] = (kb[i] + splits[i] - 1) / splits[i]; }
  for (int i = 1; i < n; ++i)
    for (int j = i; j > 0 && kbs[ord[j]] > kbs[ord[j - 1]]; --j) { const int t = ord[j]; ord[j] = ord[j - 1]; ord[j - 1] = t; }
  static thread_local std::vector<double> load;
  load.assign((size_t)P, 0.0);
  const double TILE_FIXED = 1.5;
  long t = 0;
  for (int oi = 0; oi < n; ++oi) {
    const int i = ord[oi];
    const int ns = (kb[i] + kbs[i] - 1) / kbs[i];
    for (int sp = 0; sp < ns; ++sp) {
      const int nkb = sp + 1 < ns ? kbs[i] : kb[i] - (ns - 1) * kbs[i];
      for (long m = 0; m < base_tiles[i]; ++m, ++t) {
        const long w = t / P, pos = t % P;
        load[(size_t)((snake && (w & 1)) ? P - 1 - pos : pos)] += nkb + TILE_FIXED;
      }
    }
  }
  double mx = 0.0;
  for (int q = 0; q < P; ++q) mx = load[(size_t)q] > mx ? load[(size_t)q] : mx;
  return mx + 2e-3 * (double)t;      // ties: fewer tiles (less reduce-add traffic)
}
void plan_splits(int n, const long* base_tiles, const int* kb, const bool* can_split, int P, int* splits) {
  static const int snake = getenv("K3M_GEMM_SNAKE") ? atoi(getenv("K3M_GEMM_SNAKE")) : 1;
  int kmax = 0;
  bool any = false;
  for (int i = 0; i < n; ++i) { splits[i] = 1; if (can_split[i]) { any = true; kmax = kb[i] > kmax ? kb[i] : kmax; } }
  if (!any) return;
  double best = model_makespan(n, base_tiles, kb, splits, P, snake != 0);
  int cand[8];
  for (int L = 4; L < kmax; ++L) {
    for (int i = 0; i < n; ++i) {
      int sp = 1;
      if (can_split[i]) {
        sp = (kb[i] + L / 2) / L;
        if (sp > kb[i] / 4) sp = kb[i] / 4;
        if (sp < 1) sp = 1;
      }
      cand[i] = sp;
    }
    const double m = model_makespan(n, base_tiles, kb, cand, P, snake != 0);
    if (m < best) { best = m; for (int i = 0; i < n; ++i) splits[i] = cand[i]; }
  }
}

// what the last k3m_gemm_bf16 / k3m_gemm_grouped_bf16 call of this thread dispatched (tests assert that the
// benchmark's kernel variant is the one they exercised)
thread_local char g_variant[160] = "";

struct Desc {      // == K3MGemmDesc of include/k3m.h
  const void* a; int lda; const void* b; int ldb; int layout; int M, N, K;
  void* out; int ldc; int out_f32; int accumulate; const void* bias; int act;
  const void* aux_in; int ld_aux_in; void* aux_out; int ld_aux_out;
  const void* addend; int ld_add; int addend_f32; const void* rng; unsigned int drop_site; float drop_p;
};

int check_desc(const Desc& d) {
  K3M_REQUIRE(d.M > 0 && d.N > 0 && d.K > 0, "gemm: empty problem M=%d N=%d K=%d", d.M, d.N, d.K);
  K3M_REQUIRE(d.layout >= 0 && d.layout <= 2, "gemm: layout must be 0 (nt), 1 (nn) or 2 (tn)");
  K3M_REQUIRE(d.lda % 8 == 0 && d.ldb % 8 == 0, "gemm: operand row strides must be multiples of 8 elements (lda=%d ldb=%d)", d.lda, d.ldb);
  K3M_REQUIRE(((uintptr_t)d.a & 15) == 0 && ((uintptr_t)d.b & 15) == 0, "gemm: operands must be 16-byte aligned");
  K3M_REQUIRE(!d.accumulate || d.out_f32, "gemm: accumulate needs an fp32 output");
  K3M_REQUIRE(!(d.act == ACT_DGELU || d.act == ACT_DTANH || d.act == ACT_MULAUX) || d.aux_in, "gemm: activation-derivative epilogue needs aux_in");
  K3M_REQUIRE(d.act >= ACT_NONE && d.act <= ACT_MULAUX, "gemm: unknown activation code %d", d.act);
  return K3M_OK;
}

// Fills one problem of a launch for the chosen variant (bn, ctas, ew; nbuf = staging buffers per epilogue warp of
// that kernel instantiation) and K split: operand / output / epilogue
// tensor maps, epilogue mode.  `dual`: the 2-stage two-CTAs-per-SM variant (ew = 2 there).
int setup_problem(const Desc& d, int bn, int ctas, int ew, int nbuf, int splits, ProblemArgs* out) {
  const bool a_mn = d.layout == 2, b_mn = d.layout != 0;
  const int M = d.M, N = d.N, K = d.K;
  const int total_kb = (K + BK - 1) / BK;
  // A: K-major -> tensor [M, K]; MN-major -> tensor [K, M]
  GemmParams& p = out->p;
  p.a3d = a_mn && get_map_mn3d(d.a, M, K, d.lda, BM / 64, &out->map_a);
  p.b3d = b_mn && bn / ctas >= 64 && get_map_mn3d(d.b, N, K, d.ldb, bn / ctas / 64, &out->map_b);
  if (!p.a3d && !(a_mn ? get_map(d.a, M, K, d.lda, 64, 64, &out->map_a) : get_map(d.a, K, M, d.lda, 64, BM, &out->map_a))) return K3M_ERR_CUDA;
  if (!p.b3d && !(b_mn ? get_map(d.b, N, K, d.ldb, 64, 64, &out->map_b) : get_map(d.b, K, N, d.ldb, 64, bn / ctas, &out->map_b))) return K3M_ERR_CUDA;
  p.M = M; p.N = N; p.K = K;
  p.kb_per_split = (total_kb + splits - 1) / splits;
  p.splits = (total_kb + p.kb_per_split - 1) / p.kb_per_split;
  p.tiles_m = p.tiles_n = 0;
  out->total_kb = total_kb;
  out->tile_begin = 0;
  p.out = d.out; p.ldc = d.ldc; p.out_f32 = d.out_f32; p.accumulate = d.accumulate;
  p.bias = reinterpret_cast<const bf16*>(d.bias);
  p.act = d.act;
  p.aux_in = reinterpret_cast<const bf16*>(d.aux_in); p.ld_aux_in = d.ld_aux_in;
  p.aux_out = reinterpret_cast<bf16*>(d.aux_out); p.ld_aux_out = d.ld_aux_out;
  p.addend = d.addend; p.ld_add = d.ld_add; p.addend_f32 = d.addend_f32;
  p.drop.rng = reinterpret_cast<const unsigned long long*>(d.rng);
  p.drop.site = d.drop_site;
  p.drop.thresh = (d.rng && d.drop_p > 0.f) ? (unsigned int)(d.drop_p * 65536.0f + 0.5f) : 0u;
  p.drop.scale = 1.0f / (1.0f - d.drop_p);
  p.drop_ld = (N + 7) / 8 * 8;
  auto al = [](const void* q, int ld, int esz) { return q == nullptr || ((((uintptr_t)q) & 15) == 0 && (ld * esz) % 16 == 0); };
  p.vec_ok = al(d.out, d.ldc, d.out_f32 ? 4 : 2) && al(d.bias, 8, 2) && al(d.aux_in, d.ld_aux_in, 2) && al(d.aux_out, d.ld_aux_out, 2) &&
             al(d.addend, d.ld_add, d.addend_f32 ? 4 : 2) && (!d.out_f32 || (d.ldc % 8 == 0 && (((uintptr_t)d.out) & 31) == 0));
  // TMA-store epilogue when the outputs can be described by tensor maps (16-byte aligned base and row pitch)
  out->map_c = out->map_a; out->map_aux = out->map_a; out->map_in = out->map_a;
  const int esz = d.out_f32 ? 4 : 2;
  auto tma_able = [](const void* q, long pitch_bytes) { return q != nullptr && (((uintptr_t)q) & 15) == 0 && pitch_bytes % 16 == 0; };
  static const int no_tma_store = getenv("K3M_GEMM_NO_TMA_STORE") ? atoi(getenv("K3M_GEMM_NO_TMA_STORE")) : 0;
  const bool want_aux = (d.act == ACT_GELU || d.act == ACT_GELU_GRAD) && d.aux_out != nullptr;
  p.tma_out = !no_tma_store && tma_able(d.out, (long)d.ldc * esz) && (!want_aux || (!d.out_f32 && tma_able(d.aux_out, (long)d.ld_aux_out * 2)));
  const int cw = bn / ew;
  p.tin = 0;
  if (p.tma_out) {
    const int box = (cw * esz > 128) ? 128 / esz : cw;
    if (!get_map(d.out, N, M, d.ldc, box, 32, &out->map_c, esz)) return K3M_ERR_CUDA;
    if (want_aux && !get_map(d.aux_out, N, M, d.ld_aux_out, cw, 32, &out->map_aux, 2)) return K3M_ERR_CUDA;
    // one bf16 epilogue operand can arrive through TMA into the staging buffer (same box as the output)
    static const int no_tma_in = getenv("K3M_GEMM_NO_TMA_IN") ? atoi(getenv("K3M_GEMM_NO_TMA_IN")) : 0;
    if (!no_tma_in && !d.out_f32 && !want_aux) {
      const void* src = nullptr;
      int ld_src = 0;
      if ((d.act == ACT_DGELU || d.act == ACT_DTANH || d.act == ACT_MULAUX) && tma_able(d.aux_in, (long)d.ld_aux_in * 2)) {
        p.tin = 1; src = d.aux_in; ld_src = d.ld_aux_in;
      } else if (d.addend != nullptr && !d.addend_f32 && tma_able(d.addend, (long)d.ld_add * 2)) {
        p.tin = 2; src = d.addend; ld_src = d.ld_add;
      }
      if (p.tin && !get_map(src, N, M, ld_src, cw, 32, &out->map_in, 2)) return K3M_ERR_CUDA;
    }
  }
  static const int probe = getenv("K3M_GEMM_PROBE") ? atoi(getenv("K3M_GEMM_PROBE")) : 0;
  p.probe = probe;
  p.fast = F_NONE;
  static const int no_fast = getenv("K3M_GEMM_NO_FAST") ? atoi(getenv("K3M_GEMM_NO_FAST")) : 0;
  if (!no_fast && p.tma_out && !d.out_f32 && !d.accumulate && p.vec_ok && N % 8 == 0) {
    const bool has_drop = p.drop.thresh != 0;
    if (d.act == ACT_NONE && d.aux_out == nullptr) {
      if (d.addend == nullptr) p.fast = F_PLAIN;
      else if (p.tin == 2) p.fast = F_ADD;
    } else if ((d.act == ACT_GELU || d.act == ACT_GELU_GRAD) && want_aux && d.addend == nullptr && !has_drop) {
      p.fast = F_GELU;
    } else if ((d.act == ACT_DGELU || d.act == ACT_MULAUX) && p.tin == 1 && d.addend == nullptr && d.bias == nullptr && !has_drop) {
      p.fast = F_DGELU;
    }
  }
  static const int no_facc = getenv("K3M_GEMM_NO_FACC") ? atoi(getenv("K3M_GEMM_NO_FACC")) : 0;
  if (!no_fast && !no_facc && p.tma_out && d.out_f32 && d.accumulate && d.act == ACT_NONE && d.bias == nullptr &&
      d.addend == nullptr && d.aux_out == nullptr && p.drop.thresh == 0 && N % 4 == 0) {
    p.fast = F_ACC;               // 16-column fp32 boxes (64-byte rows) for the half-buffer reduce-add epilogue
    if (!get_map(d.out, N, M, d.ldc, 16, 32, &out->map_c, 4)) return K3M_ERR_CUDA;
  }
  // F_GELU with ONE staging buffer per epilogue warp stores 32-column slabs from the two halves of that buffer
  // (fast_tile): output and saved-operand maps with 32 x 32 boxes (64-byte rows, SWIZZLE_64B)
  if (p.fast == F_GELU && nbuf == 1 && cw == 64) {
    if (!get_map(d.out, N, M, d.ldc, 32, 32, &out->map_c, 2)) return K3M_ERR_CUDA;
    if (!get_map(d.aux_out, N, M, d.ld_aux_out, 32, 32, &out->map_aux, 2)) return K3M_ERR_CUDA;
  }
  return K3M_OK;
}

const char* fast_name(int f) {
  static const char* names[] = {"generic", "F_PLAIN", "F_ADD", "F_GELU", "F_DGELU", "F_ACC"};
  return names[f];
}

}  // namespace

extern "C" const char* k3m_gemm_last_variant(void) { return g_variant; }
extern "C" void k3m_gemm_set_sm_limit(int n) { g_sm_limit = n; }

extern "C" int k3m_gemm_bf16(const void* a, int lda, const void* b, int ldb, int layout, int M, int N, int K,
                             void* out, int ldc, int out_f32, int accumulate, const void* bias, int act,
                             const void* aux_in, int ld_aux_in, void* aux_out, int ld_aux_out, const void* addend,
                             int ld_add, int addend_f32, const void* rng, unsigned int drop_site, float drop_p,
                             void* stream) {
  Desc d = {a, lda, b, ldb, layout, M, N, K, out, ldc, out_f32, accumulate, bias, act, aux_in, ld_aux_in, aux_out,
            ld_aux_out, addend, ld_add, addend_f32, rng, drop_site, drop_p};
  if (int rc = check_desc(d)) return rc;
  static const int force_bn = getenv("K3M_GEMM_BN") ? atoi(getenv("K3M_GEMM_BN")) : 0;
  int bn = pick_bn(M, N, K, force_bn);
  const int total_kb = (K + BK - 1) / BK;
  // CTA pairs (cta_group::2, 256 x 256 tile per pair): each CTA streams 32 KB per k-block instead of
  // 48 KB for the same MMA work, so the 5-stage ring covers ~1.3 us of load latency instead of ~0.8 us —
  // these K <= 3072 GEMMs are bound by bytes in flight, not by the tensor pipe.  Used when the problem
  // still fills the 74 pairs.
  static const int mode_2cta = getenv("K3M_GEMM_2CTA") ? atoi(getenv("K3M_GEMM_2CTA")) : 1;
  int ctas = 1;
  if (mode_2cta && N >= 256 && M >= 256 && force_bn == 0) {
    const long pt = (long)((M + 255) / 256) * ((N + 255) / 256);
    const long pairs = num_sms() / 2;
    if (mode_2cta == 2 || (accumulate ? (pt * 4 >= pairs && total_kb >= 32) : (pt * 5 >= pairs * 3))) {
      ctas = 2;
      bn = 256;
      // wave quantisation: 108 pair tiles on 74 pairs are 2 waves at 73 % (M = 9216 x N = 768, M = 2304 x
      // N = 3072), 76 tiles (M = 4736 x N = 1024) 2 waves at 51 %.  256 x 128 pair tiles (24 KB per k-block
      // and CTA, 6 stages) halve the granule — and were measured SLOWER on every such shape (9216x768x3072:
      // 40.5 -> 53.4 us; 2304x3072x768: 17.5 -> 19.8 us; step 22.2 -> 23.1 ms): they move 1.5x the operand
      // bytes per flop, and bytes in flight, not the wave count, is what bounds these GEMMs.  Kept behind
      // K3M_GEMM_P128=1 (rule) / 2 (always) as the record of that experiment; off by default.
      static const int mode_p128 = getenv("K3M_GEMM_P128") ? atoi(getenv("K3M_GEMM_P128")) : 0;
      if (mode_p128 && !accumulate) {
        const long pt2 = (long)((M + 255) / 256) * ((N + 127) / 128);
        const double e256 = (double)pt / (double)(pairs * ((pt + pairs - 1) / pairs));
        const double e128 = 0.92 * (double)pt2 / (double)(pairs * ((pt2 + pairs - 1) / pairs));
        if (mode_p128 == 2 || e128 > e256 + 0.04) bn = 128;
      }
    }
  }
  int splits = 1;
  static const int split_model = getenv("K3M_GEMM_SPLIT_MODEL") ? atoi(getenv("K3M_GEMM_SPLIT_MODEL")) : 1;
  if (accumulate) {
    const int tiles = ((M + BM * ctas - 1) / (BM * ctas)) * ((N + bn - 1) / bn);
    const int resident = num_sms() / ctas;
    if (split_model) {
      const long bt = tiles;
      const bool can = true;
      plan_splits(1, &bt, &total_kb, &can, resident, &splits);
    } else {
      while (tiles * splits * 2 <= resident && total_kb / (splits * 2) >= 4) splits *= 2;
    }
  }
  // epilogue warps: 16 (4 column groups) when the epilogue does real work per element, else 8
  static const int force_ew = getenv("K3M_GEMM_EW") ? atoi(getenv("K3M_GEMM_EW")) : 0;
  const bool has_drop = rng && drop_p > 0.f;
  const bool heavy = act != ACT_NONE || has_drop || addend != nullptr;
  int ew = (bn == 256 || ctas == 2) ? 4 : bn == 64 ? 1 : (force_ew ? force_ew : (heavy ? 4 : 2));
  // dual variant (K3M_GEMM_DUAL=1): single-wave 128-wide problems, where a CTA holds one tile and nothing of
  // its prologue / epilogue overlaps anything unless another CTA shares the SM.  Measured: each launch gets
  // slower alone (2304x768x768 8.4 -> 9.5 us; the step's GEMMs back to back 19.1 -> 20.5 ms) and the graph
  // step, where launches of three streams do share SMs, stays at 21.7-21.8 ms: what the sharing wins the
  // 2-stage ring loses.  Off by default.
  static const int mode_dual = getenv("K3M_GEMM_DUAL") ? atoi(getenv("K3M_GEMM_DUAL")) : 0;
  bool dual = false;
  if (mode_dual && bn == 128 && ctas == 1) {
    const int kbps = (total_kb + splits - 1) / splits;
    const long tiles = (long)((M + BM - 1) / BM) * ((N + 127) / 128) * ((total_kb + kbps - 1) / kbps);
    dual = mode_dual == 2 || tiles <= num_sms();
  }
  if (dual) ew = 2;
  GroupArgs<1> g;
  g.n = 1;
  if (int rc = setup_problem(d, bn, ctas, ew, (ew <= 2 && !dual) ? 2 : 1, splits, &g.prob[0])) return rc;
  const GemmParams& p = g.prob[0].p;
  snprintf(g_variant, sizeof(g_variant), "bn=%d ctas=%d ew=%d dual=%d splits=%d epi=%s tin=%d tma_out=%d acc=%d group=1 mn3d=%d%d",
           bn, ctas, ew, dual ? 1 : 0, p.splits, fast_name(p.fast), p.tin, p.tma_out, accumulate, p.a3d, p.b3d);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (layout == 0) return dispatch<false, false>(bn, dual ? 0 : ew, ctas, g, st);
  if (layout == 1) return dispatch<false, true>(bn, dual ? 0 : ew, ctas, g, st);
  return dispatch<true, true>(bn, dual ? 0 : ew, ctas, g, st);
}

// Several independent problems of ONE layout in one persistent launch (see GroupArgs).  Problems are ordered
// longest-k first in the tile list so that the static round-robin walk ends on the short tiles.
extern "C" int k3m_gemm_grouped_bf16(const void* descs, int n, void* stream) {
  const Desc* d = reinterpret_cast<const Desc*>(descs);
  K3M_REQUIRE(n >= 1 && n <= 8, "gemm_grouped: 1..8 problems per launch (got %d)", n);
  for (int i = 0; i < n; ++i) {
    if (int rc = check_desc(d[i])) return rc;
    K3M_REQUIRE(d[i].layout == d[0].layout, "gemm_grouped: every problem of a group must have the same layout");
    // (fp32 STORE outputs were rejected here for a while: the generic epilogue of the 128 x 128 grouped variant added
    // garbage for bias elements 3, 5, 7 of every 8 — see ld8_bits above; fixed, tests/test_kernels_gpu.py covers it)
  }
  // variant: CTA pairs (256 x 256 tile per pair) when every problem is at least one pair tile in both directions
  // and the group as a whole fills the pairs; else 128 x 128 single-CTA tiles
  static const int mode_2cta = getenv("K3M_GEMM_2CTA") ? atoi(getenv("K3M_GEMM_2CTA")) : 1;
  bool all_big = mode_2cta != 0;
  long pt = 0;
  for (int i = 0; i < n; ++i) {
    all_big = all_big && d[i].M >= 256 && d[i].N >= 256;
    pt += (long)((d[i].M + 255) / 256) * ((d[i].N + 255) / 256);
  }
  const long pairs = num_sms() / 2;
  const int ctas = (all_big && pt * 5 >= pairs * 3) ? 2 : 1;
  const int bn = ctas == 2 ? 256 : 128, ew = 4;
  // split-K of accumulating problems (weight gradients): until the group has ~3 waves of tiles
  int splits[8], order[8];
  const long resident = num_sms() / ctas;
  long tiles_total = 0;
  long base_tiles[8];
  for (int i = 0; i < n; ++i) {
    base_tiles[i] = (long)((d[i].M + BM * ctas - 1) / (BM * ctas)) * ((d[i].N + bn - 1) / bn);
    tiles_total += base_tiles[i];
    splits[i] = 1;
    order[i] = i;
  }
  static const int split_model = getenv("K3M_GEMM_SPLIT_MODEL") ? atoi(getenv("K3M_GEMM_SPLIT_MODEL")) : 1;
  if (split_model) {
    int kbv[8];
    bool can[8];
    for (int i = 0; i < n; ++i) { kbv[i] = (d[i].K + BK - 1) / BK; can[i] = d[i].accumulate != 0; }
    plan_splits(n, base_tiles, kbv, can, (int)resident, splits);
  }
  for (bool grew = !split_model; grew && tiles_total < 3 * resident;) {
    grew = false;
    for (int i = 0; i < n && tiles_total < 3 * resident; ++i) {
      const int kb = (d[i].K + BK - 1) / BK;
      if (d[i].accumulate && kb / (splits[i] * 2) >= 4) {
        tiles_total += base_tiles[i] * splits[i];
        splits[i] *= 2;
        grew = true;
      }
    }
  }
  // longest tiles first
  auto kb_of = [&](int i) { const int kb = (d[i].K + BK - 1) / BK; return (kb + splits[i] - 1) / splits[i]; };
  for (int i = 1; i < n; ++i)
    for (int j = i; j > 0 && kb_of(order[j]) > kb_of(order[j - 1]); --j) { const int t = order[j]; order[j] = order[j - 1]; order[j - 1] = t; }
  GroupArgs<8> g;
  g.n = n;
  for (int i = 0; i < n; ++i)
    if (int rc = setup_problem(d[order[i]], bn, ctas, ew, 1, splits[order[i]], &g.prob[i])) return rc;
  for (int i = n; i < 8; ++i) g.prob[i] = g.prob[0];
  snprintf(g_variant, sizeof(g_variant), "bn=%d ctas=%d ew=%d dual=0 splits=%d epi=%s tin=%d tma_out=%d acc=%d group=%d mn3d=%d%d",
           bn, ctas, ew, g.prob[0].p.splits, fast_name(g.prob[0].p.fast), g.prob[0].p.tin, g.prob[0].p.tma_out,
           d[0].accumulate, n, g.prob[0].p.a3d, g.prob[0].p.b3d);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (d[0].layout == 0) return dispatch_group<false, false>(ctas, g, st);
  if (d[0].layout == 1) return dispatch_group<false, true>(ctas, g, st);
  return dispatch_group<true, true>(ctas, g, st);
}
