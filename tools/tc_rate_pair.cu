// Bring-up probe 5 (not part of the library): issue-to-completion rate of kind::tf32 MMAs with the no-swizzle K-major
// operand layout the kernels use — one CTA (M = 128) and a CTA pair (M = 256, cta_group::2), SS and TS forms,
// N = 128 and 112.  Prints cycles per MMA for a chain of 1024 back-to-back MMAs on resident operands.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/_bin/tc_rate_pair tools/tc_rate_pair.cu && tools/_bin/tc_rate_pair
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
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
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  long long t0 = clock64();
  while (!done) {
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (clock64() - t0 > 2000000000LL) { printf("timeout\n"); asm volatile("trap;"); }
  }
}
// mode: 0 = SS, 1 = TS; PAIR = 1 -> cta_group::2.  A: 32 k-steps resident (128 KB would not fit: 8 k-steps, reused)
template <int PAIR>
__global__ void __launch_bounds__(128) rate(long long* out, int N, int mode, int reps) {
  extern __shared__ __align__(1024) uint8_t smem[];
  float* A = reinterpret_cast<float*>(smem);                    // [16 slabs][128][4] = 32 KB (8 k-steps)
  float* Bh = reinterpret_cast<float*>(smem + 32768);           // [16 slabs][rows][4]
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  uint32_t rank = 0;
  if (PAIR) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const int rows = PAIR ? N / 2 : N;
  if (warp == 0) {
    if (PAIR) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = tid; i < 16 * 128 * 4; i += 128) A[i] = 1.f;
  for (int i = tid; i < 16 * rows * 4; i += 128) Bh[i] = 0.5f;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  if (PAIR) {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  } else {
    __syncthreads();
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
  if (rank == 0 && warp == 0) {
    const uint32_t idesc = make_idesc_tf32(PAIR ? 256 : 128, N);
    const uint32_t a0 = smem_u32(A), b0 = smem_u32(Bh);
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        const uint64_t ad = make_desc(a0 + ks * 2 * 128 * 16, 128 * 16, 128);
        const uint64_t bd = make_desc(b0 + ks * 2 * rows * 16, rows * 16, 128);
        const uint32_t acc = (r | ks) ? 1u : 0u;
        if (PAIR) {
          if (mode == 0)
            asm volatile("{\n .reg .pred e, p;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %4, 0;\n"
                         " @e tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n}\n"
                         ::"r"(tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(acc) : "memory");
          else
            asm volatile("{\n .reg .pred e, p;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %4, 0;\n"
                         " @e tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], %2, %3, p;\n}\n"
                         ::"r"(tmem), "r"(tmem + 256 + ks * 8), "l"(bd), "r"(idesc), "r"(acc) : "memory");
        } else {
          if (mode == 0)
            asm volatile("{\n .reg .pred e, p;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %4, 0;\n"
                         " @e tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n"
                         ::"r"(tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(acc) : "memory");
          else
            asm volatile("{\n .reg .pred e, p;\n elect.sync _|e, 0xffffffff;\n setp.ne.b32 p, %4, 0;\n"
                         " @e tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n}\n"
                         ::"r"(tmem), "r"(tmem + 256 + ks * 8), "l"(bd), "r"(idesc), "r"(acc) : "memory");
        }
      }
    }
    const long long t1 = clock64();
    if (PAIR)
      asm volatile("{\n .reg .pred e;\n elect.sync _|e, 0xffffffff;\n"
                   " @e tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n}\n"
                   ::"r"(smem_u32(&bar)), "h"((uint16_t)3) : "memory");
    else
      asm volatile("{\n .reg .pred e;\n elect.sync _|e, 0xffffffff;\n"
                   " @e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n}\n"
                   ::"r"(smem_u32(&bar)) : "memory");
    mbar_wait(smem_u32(&bar), 0);
    const long long t2 = clock64();
    if (tid == 0 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  } else {
    mbar_wait(smem_u32(&bar), 0);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  if (PAIR) {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
  } else {
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
  }
}
int main() {
  long long* d;
  cudaMalloc(&d, 16);
  const int reps = 128;                      // 1024 MMAs
  const size_t smem = 32768 + 16 * 256 * 16;
  cudaFuncSetAttribute(rate<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(rate<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int pair = 0; pair < 2; ++pair)
    for (int mode = 0; mode < 2; ++mode)
      for (int N : {128, 112, 256}) {
        for (int grid : {1, 148}) {
          const int g = pair ? (grid == 1 ? 2 : 148) : grid;
          if (pair) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(g); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            cudaLaunchKernelEx(&cfg, rate<1>, d, N, mode, reps);
          } else {
            rate<0><<<g, 128, smem>>>(d, N, mode, reps);
          }
          long long h[2];
          cudaError_t e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
          if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 2; }
          printf("%s %s N=%3d grid=%3d: issue %.1f cycles/MMA, completion %.1f cycles/MMA\n", pair ? "pair M=256" : "single M=128",
                 mode ? "TS" : "SS", N, g, (double)h[0] / (reps * 8), (double)h[1] / (reps * 8));
        }
      }
  return 0;
}
