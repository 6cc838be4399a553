// Bring-up probe 3 (not part of the library): register layout of tcgen05.ld.16x256b.x2 against a TMEM tile written with
// the 32x32b shape (lane = row, column = word).  Prints, for threads 0..7 and 28..31, which (row, col) each register holds.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/_bin/tc_probe3 tools/tc_probe3.cu && tools/_bin/tc_probe3
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void __launch_bounds__(128) probe(uint32_t* out) {
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(32));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
  uint32_t v[16];
  for (int i = 0; i < 16; ++i) v[i] = (uint32_t)((warp * 32 + lane) * 100 + i);      // value = row * 100 + col
  const uint32_t addr = tmem + ((uint32_t)(warp * 32) << 16);
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
               ::"r"(addr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                 "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int half = 0; half < 2; ++half) {
    uint32_t r[8];
    const uint32_t a2 = tmem + ((uint32_t)(warp * 32 + half * 16) << 16);
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(a2));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 8; ++i) out[((warp * 2 + half) * 32 + lane) * 8 + i] = r[i];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(32));
}
int main() {
  uint32_t* d;
  cudaMalloc(&d, 4 * 2 * 32 * 8 * 4);
  probe<<<1, 128>>>(d);
  uint32_t h[4 * 2 * 32 * 8];
  cudaError_t e = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 2; }
  for (int w = 0; w < 2; ++w)
    for (int half = 0; half < 2; ++half)
      for (int t = 0; t < 32; ++t) {
        if (t >= 8 && t < 28) continue;
        printf("warp %d half %d thread %2d:", w, half, t);
        for (int i = 0; i < 8; ++i) { const uint32_t x = h[((w * 2 + half) * 32 + t) * 8 + i]; printf(" r%d=(%u,%u)", i, x / 100, x % 100); }
        printf("\n");
      }
  return 0;
}
