// Bring-up probe 4 (not part of the library): tcgen05.mma.cta_group::2 with kind::tf32 on a CTA pair.
// Checks, against integers computed on the host,
//   * how the operands split over the pair: A = 128 rows per CTA (own shared memory / own TMEM), B = N/2 rows per CTA,
//     and WHICH half of the N columns each CTA's rows supply,
//   * N = 128 and N = 112 (56 rows per CTA) with M = 256,
//   * the .ts form (A from each CTA's tensor memory),
//   * tcgen05.commit ... multicast::cluster onto a barrier in both CTAs, a remote mbarrier arrive (mapa) from the
//     peer onto the leader's barrier, and tcgen05.alloc / dealloc with cta_group::2.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/_bin/tc_probe_pair tools/tc_probe_pair.cu && tools/_bin/tc_probe_pair
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
__device__ __forceinline__ uint32_t make_idesc_tf32(int m, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  long long t0 = clock64();
  while (!done) {
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (clock64() - t0 > 2000000000LL) { printf("timeout bar %u\n", bar); asm volatile("trap;"); }
  }
}
__device__ int a_val(int g, int k) { return ((g * 3 + k * 7) % 11) - 5; }
__device__ int b_val(int n, int k) { return ((n * 5 + k * 3) % 13) - 6; }

// out[test][256][128]; test 0 = SS, test 1 = TS
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128) probe(float* out, int N) {
  __shared__ __align__(1024) float A[2 * 128 * 4];
  __shared__ __align__(1024) float Bh[2 * 64 * 4];
  __shared__ __align__(8) uint64_t bars[3];
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const int NH = N / 2;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
  }
  if (tid == 0) {
    for (int i = 0; i < 3; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bars[i])), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // operands: A rows of this CTA (K-major, no swizzle: [2 slabs][128 rows][4]), B rows rank*NH .. rank*NH+NH-1
  const int g = (int)rank * 128 + tid;
  for (int k = 0; k < 8; ++k) A[((k >> 2) * 128 + tid) * 4 + (k & 3)] = (float)a_val(g, k);
  if (tid < NH)
    for (int k = 0; k < 8; ++k) Bh[((k >> 2) * NH + tid) * 4 + (k & 3)] = (float)b_val((int)rank * NH + tid, k);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
  // A also into tensor memory (columns 256..263) for the .ts test: lane = row
  {
    uint32_t v[8];
    for (int k = 0; k < 8; ++k) v[k] = __float_as_uint((float)a_val(g, k));
    const uint32_t addr = tmem + 256 + ((uint32_t)(warp * 32) << 16);
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"r"(addr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t bar_done = smem_u32(&bars[0]), bar_peer = smem_u32(&bars[1]);
  if (rank == 1 && tid == 0) {
    // the peer tells the leader that its operands are in place: remote arrive on the LEADER's barrier
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(bar_peer), "r"(0));
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
  }
  if (rank == 0 && tid == 0) {
    mbar_wait_cluster(bar_peer, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t idesc = make_idesc_tf32(256, N);
    const uint64_t ad = make_desc(smem_u32(A), 128 * 16, 128), bd = make_desc(smem_u32(Bh), NH * 16, 128);
    asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
                 " tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n}\n"
                 ::"r"(tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(0) : "memory");
    asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
                 " tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], %2, %3, p;\n}\n"
                 ::"r"(tmem + 128), "r"(tmem + 256), "l"(bd), "r"(idesc), "r"(0) : "memory");
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar_done), "h"((uint16_t)3) : "memory");
  }
  mbar_wait_cluster(bar_done, 0);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int test = 0; test < 2; ++test)
    for (int c0 = 0; c0 < N; c0 += 8) {
      uint32_t r[8];
      const uint32_t addr = tmem + test * 128 + c0 + ((uint32_t)(warp * 32) << 16);
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(addr));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      for (int i = 0; i < 8; ++i) out[((size_t)test * 256 + g) * 128 + c0 + i] = __uint_as_float(r[i]);
    }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

static int a_h(int g, int k) { return ((g * 3 + k * 7) % 11) - 5; }
static int b_h(int n, int k) { return ((n * 5 + k * 3) % 13) - 6; }
int main() {
  float* d;
  const size_t n = 2 * 256 * 128;
  cudaMalloc(&d, n * 4);
  float* h = (float*)malloc(n * 4);
  const int Ns[2] = {128, 112};
  for (int t = 0; t < 2; ++t) {
    const int N = Ns[t];
    cudaMemset(d, 0xff, n * 4);
    probe<<<2, 128>>>(d, N);
    cudaError_t e = cudaMemcpy(h, d, n * 4, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { printf("N=%d CUDA error %s\n", N, cudaGetErrorString(e)); return 2; }
    for (int test = 0; test < 2; ++test) {
      long bad_same = 0, bad_swapped = 0;
      for (int g = 0; g < 256; ++g)
        for (int c = 0; c < N; ++c) {
          int ref = 0, refs = 0;
          const int cs = (c + N / 2) % N;                            // column if the halves were swapped
          for (int k = 0; k < 8; ++k) { ref += a_h(g, k) * b_h(c, k); refs += a_h(g, k) * b_h(cs, k); }
          const float v = h[((size_t)test * 256 + g) * 128 + c];
          bad_same += v != (float)ref;
          bad_swapped += v != (float)refs;
        }
      printf("N=%d %s: mismatches with leader = columns [0,N/2): %ld, with the halves swapped: %ld  (D[0][0..3] = %g %g %g %g, D[128][0..1] = %g %g)\n",
             N, test ? "TS" : "SS", bad_same, bad_swapped, h[(size_t)test * 256 * 128], h[(size_t)test * 256 * 128 + 1],
             h[(size_t)test * 256 * 128 + 2], h[(size_t)test * 256 * 128 + 3], h[((size_t)test * 256 + 128) * 128],
             h[((size_t)test * 256 + 128) * 128 + 1]);
    }
  }
  return 0;
}
