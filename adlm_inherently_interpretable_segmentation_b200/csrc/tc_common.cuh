// tcgen05 / TMEM / mbarrier primitives used by the tensor-core kernels (inline PTX, sm_100a).
//
// Conventions verified on a B200 with tools/tc_probe.cu (profiles/r1_tcgen05_probe.log):
//   * shared-memory operands use the no-swizzle K-major canonical layout: an [R x K] fp32/tf32 operand is
//     stored as [K/4 slabs][R rows][4 k] (one 16-byte row per (row, k-quad); a core matrix = 8 rows x 16 B
//     = 128 contiguous bytes).  Descriptor: LBO = byte distance between the two k-quads of one MMA
//     (= R*16), SBO = byte distance between 8-row groups (= 128).  One kind::tf32 MMA consumes K = 8,
//     i.e. two slabs; the start address advances by 2*R*16 per k-step.
//   * an A operand in TMEM holds row m in lane m and k in consecutive 32-bit columns (".ts" form).
//   * the 128-byte-swizzled K-major layout also works (tools/tc_probe2.cu, profiles/r1_tcgen05_probe2.log):
//     rows of 128 B (32 floats), 8-row atoms of 1024 B, 16-byte chunk index XOR (row % 8); SBO = 1024,
//     a k-step of 8 floats advances the start address by 32 B inside the atom, the next 32-float k-atom
//     lies rows*128 B further.  It is used where one warp writes a COLUMN of the operand (conflict-free).
//   * MN-major tf32 operands (no swizzle or SWIZZLE_128B) return zeros on this part; nothing here uses
//     them — operands needed in both orientations are stored twice.
//   * every MMA here uses M = 128 (row r of the accumulator = TMEM lane r).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace adlm {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;   // descriptor version 1 (sm_100)
  return d;                 // layout type 0: no swizzle
}

__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;            // LBO unused for swizzled K-major
  d |= (uint64_t)(1024 >> 4) << 32;  // SBO: 8-row group
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;            // SWIZZLE_128B
  return d;
}
// float offset of element (r, c) in a [rows x cols] fp32 tile stored with the 128-byte swizzle
__device__ __forceinline__ int off_sw128(int r, int c, int rows) {
  const int ci = c & 31;
  return (c >> 5) * rows * 32 + (r >> 3) * 256 + (r & 7) * 32 + (((ci >> 2) ^ (r & 7)) << 2) + (ci & 3);
}

// kind::tf32, fp32 accumulate, both operands K-major
__device__ __forceinline__ uint32_t make_idesc_tf32(int m, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ void mma_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
               " tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n"
               ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
               " tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n}\n"
               ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrives on the mbarrier when every MMA issued so far by this thread has completed
__device__ __forceinline__ void mma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// Whole-warp forms: every lane of a CONVERGED warp calls them with warp-uniform arguments and one elected lane issues.
// With the issue inside `if (lane == 0)` the compiler treats every operand as per-thread data and wraps each
// tcgen05.mma in an elect / R2UR / branch loop to move the descriptors into uniform registers (~40 SASS instructions,
// 80-130 cycles per MMA in the phase clocks); with uniform control flow around elect.sync the descriptors are computed
// in the uniform datapath.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n .reg .pred p;\n elect.sync _|p, 0xffffffff;\n selp.u32 %0, 1, 0, p;\n}\n" : "=r"(pred));
  return pred != 0u;
}
// elect.sync and the predicated instruction sit in ONE asm block: with `if (elect_one()) mma(...)` the compiler may
// sink the descriptor arithmetic into the conditional and emit a divergent region (BSSY / BRA / BSYNC.RECONVERGENT
// per MMA: ~160 cycles per instruction in the large-head forward's K loop, ncu source page, round 2); this form
// leaves it nothing to branch around.
__device__ __forceinline__ void mma_ss_w(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n .reg .pred e, p;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %4, 0;\n"
               " @e tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n"
               ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_ts_w(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n .reg .pred e, p;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %4, 0;\n"
               " @e tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n}\n"
               ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_commit_w(uint32_t bar) {
  asm volatile("{\n .reg .pred e;\n elect.sync _|e, 0xffffffff;\n"
               " @e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n}\n" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy shared-memory writes -> visible to the async proxy (UMMA operand reads)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                 "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
               ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                 "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// Bounded wait: a pipeline bug must not hang the GPU box — trap after ~2 s instead.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
               : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  if (ok) return;
  const long long t0 = clock64();
  while (true) {
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) return;
    if (clock64() - t0 > 4000000000LL) __trap();
  }
}
// one non-blocking look at a barrier phase
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n .reg .pred p;\n mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
               : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0u;
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ void named_bar_arrive(int id, int nthreads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ float lg2_approx(float x) {
  float r;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// ---- CTA-pair forms (cluster of two CTAs, tcgen05 ... cta_group::2; conventions verified with tools/tc_probe_pair.cu,
// profiles/r2_probe_pair.log): one MMA has M = 256 — accumulator rows 0..127 land in the leader's tensor memory, rows
// 128..255 in the peer's — each CTA supplies ITS 128 rows of A (own shared memory at the descriptor's address, or own
// tensor memory for .ts) and N/2 rows of B: the leader (cluster rank 0) the accumulator columns [0, N/2), the peer
// [N/2, N).  Only the leader issues; tcgen05.commit ... multicast::cluster arrives on the barrier at the same
// shared-memory offset in both CTAs.
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t dst_smem, uint32_t ncols) {      // one warp in EACH CTA
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void mma_pair_ss_w(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n .reg .pred e, p;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %4, 0;\n"
               " @e tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n}\n"
               ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_pair_ts_w(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n .reg .pred e, p;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %4, 0;\n"
               " @e tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], %2, %3, p;\n}\n"
               ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// The same with the descriptors given as (low word, high word): consecutive k-steps of an operand differ only in the
// start-address field, so the caller steps the low word by (bytes >> 4) instead of rebuilding a descriptor per MMA
// (make_desc costs four integer instructions per operand, and the issuing warp has < 87 cycles per MMA to spend).
__device__ __forceinline__ void mma_pair_ss_lh(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t a_hi, uint32_t b_hi,
                                               uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n .reg .pred e, p;\n .reg .b64 da, db;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %6, 0;\n"
               " mov.b64 da, {%1, %3};\n mov.b64 db, {%2, %4};\n"
               " @e tcgen05.mma.cta_group::2.kind::tf32 [%0], da, db, %5, p;\n}\n"
               ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(a_hi), "r"(b_hi), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_pair_ts_lh(uint32_t tmem_d, uint32_t tmem_a, uint32_t b_lo, uint32_t b_hi, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile("{\n .reg .pred e, p;\n .reg .b64 db;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %5, 0;\n"
               " mov.b64 db, {%2, %3};\n"
               " @e tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], db, %4, p;\n}\n"
               ::"r"(tmem_d), "r"(tmem_a), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate) : "memory");
}
// arrives on `bar` (same offset) in both CTAs when every MMA issued so far by this thread has completed
__device__ __forceinline__ void mma_pair_commit_w(uint32_t bar) {
  asm volatile("{\n .reg .pred e;\n elect.sync _|e, 0xffffffff;\n"
               " @e tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n}\n"
               ::"r"(bar), "h"((uint16_t)3) : "memory");
}
// arrive on the barrier at this CTA's offset `bar` in CTA `cta` of the cluster.  Default semantics (release at CTA
// scope), as CUTLASS's ClusterBarrier::arrive does: with .release.cluster ptxas puts MEMBAR.ALL.GPU in front of every
// arrive (~1000 cycles; it serialised the large-head forward's chunk pipeline, profiles/r2_cfg5_experiments.log).  What a
// remote arrival publishes here is either nothing (a relayed TMA completion) or data its own warp wrote and fenced
// (fence.proxy.async / tcgen05.fence::before_thread_sync) before the arrive.  Waiters use the plain mbar_wait.
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar, uint32_t cta) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(bar), "r"(cta));
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}

// round-to-nearest fp32 -> tf32 (the tensor core would otherwise truncate the low 13 mantissa bits)
__device__ __forceinline__ float rna_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

// Same rounding in one integer add: magnitude + half an ulp(tf32).  The 13 low bits stay in place — the tensor
// core ignores them — so this form is for MMA operands whose exact rounded value is not needed elsewhere
// (similarities, gradients).  Finite inputs only.
__device__ __forceinline__ float rna_tf32_fast(float x) { return __uint_as_float(__float_as_uint(x) + 0x1000u); }
// ... and with the low bits cleared (two instructions instead of cvt.rna's four): operands whose rounded
// value also enters fp32 arithmetic (|z|^2, |p|^2, the lo part of a 3xTF32 split)
__device__ __forceinline__ float rna_tf32_mask(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}

// Stage a row-major [rows_valid x K] fp32 matrix (prototypes) as a canonical K-major operand
// [KP/4][rows_pad][4]: value = rna(x - centre) (hi) and, for 3xTF32, lo = rna((x - centre) - hi).
// Rows >= rows_valid and columns >= K are zero.  Fast path (K == KP, 16-byte aligned base): every thread
// reads float4s, four loads in flight per thread.
template <int KP, bool X3, int BATCH = 4>
__device__ __forceinline__ void stage_rows_kmajor(float* __restrict__ hi, float* __restrict__ lo,
                                                  const float* __restrict__ src, int rows_valid, int K,
                                                  int rows_pad, float centre, int tid, int nthreads) {
  constexpr int Q = KP / 4;
  const int total = rows_pad * Q;
  if (K == KP && (reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
    // a warp covers 8 consecutive rows x 4 consecutive 4-column groups: each quarter-warp's 16-byte
    // shared-memory stores are 128 contiguous bytes (conflict-free) and the global side reads 64 contiguous
    // bytes per row (8 lines per warp load instead of 32)
    const float4* src4 = reinterpret_cast<const float4*>(src);
    const int n8 = rows_pad >> 3;                             // rows_pad is a multiple of 32, Q of 4
    for (int f0 = 0; f0 < total; f0 += nthreads * BATCH) {
      float4 v[BATCH];
#pragma unroll
      for (int u = 0; u < BATCH; ++u) {
        const int f = f0 + u * nthreads + tid;
        const int g = f >> 5, l = f & 31;                     // group of 32 = 8 rows x 4 column groups
        const int k4 = (g / n8) * 4 + (l >> 3), n = (g % n8) * 8 + (l & 7);
        v[u] = (f < total && n < rows_valid) ? __ldg(src4 + (size_t)n * Q + k4) : make_float4(centre, centre, centre, centre);
      }
#pragma unroll
      for (int u = 0; u < BATCH; ++u) {
        const int f = f0 + u * nthreads + tid;
        if (f < total) {
          const float x[4] = {v[u].x - centre, v[u].y - centre, v[u].z - centre, v[u].w - centre};
          float h[4], l[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            h[j] = rna_tf32_mask(x[j]);
            l[j] = X3 ? rna_tf32_mask(x[j] - h[j]) : 0.f;
          }
          const int g = f >> 5, li = f & 31;
          const size_t o = ((size_t)((g / n8) * 4 + (li >> 3)) * rows_pad + (g % n8) * 8 + (li & 7)) * 4;
          *reinterpret_cast<float4*>(hi + o) = make_float4(h[0], h[1], h[2], h[3]);
          if (X3) *reinterpret_cast<float4*>(lo + o) = make_float4(l[0], l[1], l[2], l[3]);
        }
      }
    }
  } else {
    for (int idx = tid; idx < total; idx += nthreads) {
      const int n = idx % rows_pad, k4 = idx / rows_pad;
      float h[4], l[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = k4 * 4 + j;
        const float x = (n < rows_valid && k < K) ? __ldg(src + (size_t)n * K + k) - centre : 0.f;
        h[j] = rna_tf32_mask(x);
        l[j] = X3 ? rna_tf32_mask(x - h[j]) : 0.f;
      }
      *reinterpret_cast<float4*>(hi + (size_t)idx * 4) = make_float4(h[0], h[1], h[2], h[3]);
      if (X3) *reinterpret_cast<float4*>(lo + (size_t)idx * 4) = make_float4(l[0], l[1], l[2], l[3]);
    }
  }
}

// Stage W [C x N] (row-major) TRANSPOSED as a K-major operand: rows = prototypes n, K index = class c,
// layout [C8/4][rows_pad][4]; value = rna(scale * W[c][n]), zero beyond C / N or when !valid.  Consecutive
// threads take consecutive n: coalesced global reads, contiguous 16-byte shared-memory stores.
__device__ __forceinline__ void stage_wt_kmajor(float* __restrict__ dst, const float* __restrict__ W, int C, int N,
                                                int C8, int rows_pad, float scale, bool valid, int tid,
                                                int nthreads) {
  const int total = (C8 >> 2) * rows_pad;
  for (int f0 = 0; f0 < total; f0 += nthreads * 4) {
    float v[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int f = f0 + u * nthreads + tid;
      const int c4 = f / rows_pad, n = f - c4 * rows_pad;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int c = c4 * 4 + j;
        v[u][j] = (valid && f < total && c < C && n < N) ? __ldg(W + (size_t)c * N + n) : 0.f;
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int f = f0 + u * nthreads + tid;
      if (f < total)
        *reinterpret_cast<float4*>(dst + (size_t)f * 4) =
            make_float4(rna_tf32_mask(scale * v[u][0]), rna_tf32_mask(scale * v[u][1]),
                        rna_tf32_mask(scale * v[u][2]), rna_tf32_mask(scale * v[u][3]));
    }
  }
}

// stage_rows_kmajor + stage_wt_kmajor with EVERY global load of both in flight before the first conversion: one
// global latency for the one-time operands instead of one per batch (the dP/dW kernel spent 17 k cycles here in six
// batches).  PB * nthreads float4 cover P, WB * nthreads quads cover W^T; the caller checks the fast-path
// conditions (K == KP, 16-byte aligned P) and the capacities, else it calls the two functions above.
template <int KP, int PB, int WB>
__device__ __forceinline__ void stage_rows_and_wt_once(float* __restrict__ pk, float* __restrict__ wt,
                                                       const float* __restrict__ P, const float* __restrict__ W,
                                                       int N, int C, int C8, int rows_pad, float centre,
                                                       bool valid_w, int tid, int nthreads) {
  constexpr int Q = KP / 4;
  const int total_p = rows_pad * Q, total_w = (C8 >> 2) * rows_pad;
  const int n8 = rows_pad >> 3;
  const float4* src4 = reinterpret_cast<const float4*>(P);
  float wv[WB][4];
  float4 pv[PB];
#pragma unroll
  for (int u = 0; u < WB; ++u) {
    const int f = u * nthreads + tid;
    const int c4 = f / rows_pad, n = f - c4 * rows_pad;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = c4 * 4 + j;
      wv[u][j] = (valid_w && f < total_w && c < C && n < N) ? __ldg(W + (size_t)c * N + n) : 0.f;
    }
  }
#pragma unroll
  for (int u = 0; u < PB; ++u) {
    const int f = u * nthreads + tid;
    const int g = f >> 5, l = f & 31;                         // group of 32 = 8 rows x 4 column groups
    const int k4 = (g / n8) * 4 + (l >> 3), n = (g % n8) * 8 + (l & 7);
    pv[u] = (f < total_p && n < N) ? __ldg(src4 + (size_t)n * Q + k4) : make_float4(centre, centre, centre, centre);
  }
#pragma unroll
  for (int u = 0; u < WB; ++u) {
    const int f = u * nthreads + tid;
    if (f < total_w)
      *reinterpret_cast<float4*>(wt + (size_t)f * 4) =
          make_float4(rna_tf32_mask(wv[u][0]), rna_tf32_mask(wv[u][1]), rna_tf32_mask(wv[u][2]), rna_tf32_mask(wv[u][3]));
  }
#pragma unroll
  for (int u = 0; u < PB; ++u) {
    const int f = u * nthreads + tid;
    if (f < total_p) {
      const int g = f >> 5, li = f & 31;
      const size_t o = ((size_t)((g / n8) * 4 + (li >> 3)) * rows_pad + (g % n8) * 8 + (li & 7)) * 4;
      *reinterpret_cast<float4*>(pk + o) =
          make_float4(rna_tf32_mask(pv[u].x - centre), rna_tf32_mask(pv[u].y - centre),
                      rna_tf32_mask(pv[u].z - centre), rna_tf32_mask(pv[u].w - centre));
    }
  }
}

// Stage W [C x N] as a K-major operand with rows = classes c, K index = prototype n: [Npad/4][Cpad][4];
// value = rna(scale * W[c][n]), zero beyond C / N.  A warp covers 8 classes x 4 prototype quads: conflict-free
// 16-byte shared-memory stores, 64 contiguous bytes per class row on the global side.
__device__ __forceinline__ void stage_w_kmajor(float* __restrict__ dst, const float* __restrict__ W, int C, int N,
                                               int Cpad, int Npad, float scale, int tid, int nthreads) {
  const int total = (Npad >> 2) * Cpad;
  const int c8 = Cpad >> 3;                                   // Cpad is a multiple of 8, Npad of 16
  for (int f0 = 0; f0 < total; f0 += nthreads * 4) {
    float v[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int f = f0 + u * nthreads + tid;
      const int g = f >> 5, l = f & 31;                       // group of 32 = 8 classes x 4 prototype quads
      const int n4 = (g / c8) * 4 + (l >> 3), c = (g % c8) * 8 + (l & 7);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = n4 * 4 + j;
        v[u][j] = (f < total && c < C && n < N) ? __ldg(W + (size_t)c * N + n) : 0.f;
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int f = f0 + u * nthreads + tid;
      const int g = f >> 5, l = f & 31;
      if (f < total)
        *reinterpret_cast<float4*>(dst + ((size_t)((g / c8) * 4 + (l >> 3)) * Cpad + (g % c8) * 8 + (l & 7)) * 4) =
            make_float4(rna_tf32_mask(scale * v[u][0]), rna_tf32_mask(scale * v[u][1]),
                        rna_tf32_mask(scale * v[u][2]), rna_tf32_mask(scale * v[u][3]));
    }
  }
}

// Stage P [N x K] TRANSPOSED (rows = depth k, K index = prototype n): [Npad/4][KP][4], same centring and
// rounding as stage_rows_kmajor.  Consecutive threads take consecutive k: coalesced reads, contiguous stores.
template <int KP>
__device__ __forceinline__ void stage_pt_kmajor(float* __restrict__ dst, const float* __restrict__ P, int N, int K,
                                                int Npad, float centre, int tid, int nthreads) {
  const int total = (Npad >> 2) * KP;
  for (int f0 = 0; f0 < total; f0 += nthreads * 4) {
    float v[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int f = f0 + u * nthreads + tid;
      const int n4 = f / KP, k = f - n4 * KP;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = n4 * 4 + j;
        v[u][j] = (f < total && n < N && k < K) ? __ldg(P + (size_t)n * K + k) : centre;
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int f = f0 + u * nthreads + tid;
      if (f < total)
        *reinterpret_cast<float4*>(dst + (size_t)f * 4) =
            make_float4(rna_tf32_mask(v[u][0] - centre), rna_tf32_mask(v[u][1] - centre),
                        rna_tf32_mask(v[u][2] - centre), rna_tf32_mask(v[u][3] - centre));
    }
  }
}

// ---- one-time operands through a raw shared-memory copy --------------------------------------------------
// cp.async puts every byte of P and W in flight at once (no registers, one global latency); the operand layouts
// are then produced by shared-to-shared transforms.  P rows get a pitch of K + 4 floats so that both "lanes over
// rows" and "lanes over columns" reads are conflict-free.  Requirements (else use the stage_* functions above):
// K == KP, P 16-byte aligned, W 8-byte aligned with an even element count.
__device__ __forceinline__ bool raw_stage_ok(const float* P, const float* W, int K, int KP, int CN) {
  return K == KP && (reinterpret_cast<uintptr_t>(P) & 15u) == 0 && (reinterpret_cast<uintptr_t>(W) & 7u) == 0 && (CN & 1) == 0;
}
__device__ __forceinline__ void raw_copy_async(float* __restrict__ praw, float* __restrict__ wraw,
                                               const float* __restrict__ P, const float* __restrict__ W, int N, int K,
                                               int CN, int tid, int nthreads) {
  const int cpr = K >> 2, pitch = K + 4;
  for (int i = tid; i < N * cpr; i += nthreads) {
    const int r = i / cpr, c = i - r * cpr;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(praw + r * pitch + c * 4)), "l"(P + (size_t)r * K + c * 4) : "memory");
  }
  for (int i = tid; i < (CN >> 1); i += nthreads)
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(wraw + i * 2)), "l"(W + i * 2) : "memory");
  asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void raw_copy_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Pk [KP/4][rows_pad][4] (rows = prototypes) from the raw copy
template <int KP>
__device__ __forceinline__ void pk_from_raw(float* __restrict__ dst, const float* __restrict__ praw, int N, int rows_pad,
                                            float centre, int tid, int nthreads) {
  const int total = (KP >> 2) * rows_pad;
  for (int f = tid; f < total; f += nthreads) {
    const int k4 = f / rows_pad, n = f - k4 * rows_pad;
    float4 v = make_float4(centre, centre, centre, centre);
    if (n < N) v = *reinterpret_cast<const float4*>(praw + n * (KP + 4) + k4 * 4);
    *reinterpret_cast<float4*>(dst + (size_t)f * 4) =
        make_float4(rna_tf32_mask(v.x - centre), rna_tf32_mask(v.y - centre), rna_tf32_mask(v.z - centre), rna_tf32_mask(v.w - centre));
  }
}
// Pt [Npad/4][KP][4] (rows = depth k, K index = prototype)
template <int KP>
__device__ __forceinline__ void pt_from_raw(float* __restrict__ dst, const float* __restrict__ praw, int N, int Npad,
                                            float centre, int tid, int nthreads) {
  const int total = (Npad >> 2) * KP;
  for (int f = tid; f < total; f += nthreads) {
    const int n4 = f / KP, k = f - n4 * KP;
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n4 * 4 + j;
      v[j] = n < N ? rna_tf32_mask(praw[n * (KP + 4) + k] - centre) : 0.f;
    }
    *reinterpret_cast<float4*>(dst + (size_t)f * 4) = make_float4(v[0], v[1], v[2], v[3]);
  }
}
// Wt [C8/4][rows_pad][4] (rows = prototypes, K index = class)
__device__ __forceinline__ void wt_from_raw(float* __restrict__ dst, const float* __restrict__ wraw, int C, int N, int C8,
                                            int rows_pad, float scale, bool valid, int tid, int nthreads) {
  const int total = (C8 >> 2) * rows_pad;
  for (int f = tid; f < total; f += nthreads) {
    const int c4 = f / rows_pad, n = f - c4 * rows_pad;
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = c4 * 4 + j;
      v[j] = (valid && c < C && n < N) ? rna_tf32_mask(scale * wraw[c * N + n]) : 0.f;
    }
    *reinterpret_cast<float4*>(dst + (size_t)f * 4) = make_float4(v[0], v[1], v[2], v[3]);
  }
}
// Wk [Npad/4][Cpad][4] (rows = classes, K index = prototype)
__device__ __forceinline__ void wk_from_raw(float* __restrict__ dst, const float* __restrict__ wraw, int C, int N, int Cpad,
                                            int Npad, float scale, int tid, int nthreads) {
  const int total = (Npad >> 2) * Cpad;
  for (int f = tid; f < total; f += nthreads) {
    const int n4 = f / Cpad, c = f - n4 * Cpad;
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n4 * 4 + j;
      v[j] = (c < C && n < N) ? rna_tf32_mask(scale * wraw[c * N + n]) : 0.f;
    }
    *reinterpret_cast<float4*>(dst + (size_t)f * 4) = make_float4(v[0], v[1], v[2], v[3]);
  }
}

// zero a shared-memory region of n4 float4s
__device__ __forceinline__ void smem_zero4(float* dst, int n4, int tid, int nthreads) {
  for (int i = tid; i < n4; i += nthreads) reinterpret_cast<float4*>(dst)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

}  // namespace tc
}  // namespace adlm
