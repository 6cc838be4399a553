// Bring-up probe for the tcgen05 path (not part of the library): checks, on a real B200, the shared-memory
// matrix-descriptor conventions this repo's kernels rely on.  One variant per process:
//   tc_probe <variant>
// Every variant computes D[128x64] = A[128x16] * B[64x16]^T with kind::tf32 from exactly representable inputs,
// so any mismatch is a layout/descriptor error.  Prints "variant V: max_err E  (first bad m,n got want)".
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 2; } } while (0)

constexpr int M = 128, N = 64, KT = 16;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version 1 (Blackwell)
  return d;                // layout_type 0 = no swizzle
}

__device__ __forceinline__ uint32_t make_idesc(int m, int n, int a_major, int b_major) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_major << 15) | ((uint32_t)b_major << 16) |
         ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {
  const long long t0 = clock64();
  while (true) {
    uint32_t ok;
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) return true;
    if (clock64() - t0 > 2000000000LL) return false;
  }
}

// K-major canonical (no swizzle): element (r, k) of an [R x KT] operand at  (k/4)*R*4 + r*4 + k%4   (floats)
//   -> core matrix = 8 rows x 16 B contiguous (128 B); next 8 rows +128 B; next 4 k  + R*16 B
__device__ __forceinline__ int off_kmajor(int r, int k, int R) { return (k >> 2) * R * 4 + r * 4 + (k & 3); }
// MN-major canonical (no swizzle): element (r, k) at (r/4)*KT*4 + k*4 + r%4
//   -> core matrix = 8 k-rows x 16 B (4 consecutive r); next 8 k +128 B; next 4 r + KT*16 B
__device__ __forceinline__ int off_mnmajor(int r, int k) { return (r >> 2) * KT * 4 + k * 4 + (r & 3); }

__global__ void __launch_bounds__(128) probe(const float* __restrict__ A, const float* __restrict__ B,
                                             float* __restrict__ D, int variant, int* __restrict__ status) {
  extern __shared__ __align__(1024) uint8_t smem[];
  float* sA = reinterpret_cast<float*>(smem);
  float* sB = sA + M * KT;
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool a_mn = (variant == 2 || variant == 3);
  const bool b_mn = (variant == 4 || variant == 5);
  const bool ts = (variant == 6);

  for (int idx = tid; idx < M * KT; idx += 128) {
    const int m = idx / KT, k = idx % KT;
    sA[a_mn ? off_mnmajor(m, k) : off_kmajor(m, k, M)] = A[idx];
  }
  for (int idx = tid; idx < N * KT; idx += 128) {
    const int n = idx / KT, k = idx % KT;
    sB[b_mn ? off_mnmajor(n, k) : off_kmajor(n, k, N)] = B[idx];
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(128));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;

  if (ts || variant == 7) {
    // A into TMEM columns [64, 80): lane = row m, column = k
    uint32_t v[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = __float_as_uint(A[(warp * 32 + lane) * KT + k]);
    const uint32_t addr = tmem + ((uint32_t)(warp * 32) << 16) + 64;
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                 ::"r"(addr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                   "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }

  if (variant == 7) {   // TMEM st/ld round trip only
    uint32_t r[16];
    const uint32_t addr = tmem + ((uint32_t)(warp * 32) << 16) + 64;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(addr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int k = 0; k < 16; ++k) D[(warp * 32 + lane) * N + k] = __uint_as_float(r[k]);
  } else {
    if (tid == 0) {
      const uint32_t idesc = make_idesc(M, N, a_mn ? 1 : 0, b_mn ? 1 : 0);
#pragma unroll
      for (int ks = 0; ks < KT / 8; ++ks) {
        // K-major: one instruction covers 2 k-slabs (8 k): start += 2 * R*16 B per step.
        //   hypothesis H0 (variants 0,2,4,6): LBO = K-direction core-matrix stride, SBO = MN-direction stride
        //   hypothesis H1 (variants 1,3,5):   swapped
        uint32_t a_lbo, a_sbo, b_lbo, b_sbo, a_start, b_start;
        if (!a_mn) { a_lbo = M * 16; a_sbo = 128; a_start = smem_u32(sA) + ks * 2 * M * 16; }
        else       { a_lbo = 128; a_sbo = KT * 16; a_start = smem_u32(sA) + ks * 128; }   // K dir: +128 B per 8 k; MN dir: +KT*16 per 4 m
        if (!b_mn) { b_lbo = N * 16; b_sbo = 128; b_start = smem_u32(sB) + ks * 2 * N * 16; }
        else       { b_lbo = 128; b_sbo = KT * 16; b_start = smem_u32(sB) + ks * 128; }
        if (variant == 1 || variant == 3) { uint32_t t = a_lbo; a_lbo = a_sbo; a_sbo = t; }
        if (variant == 1 || variant == 5) { uint32_t t = b_lbo; b_lbo = b_sbo; b_sbo = t; }
        const uint64_t adesc = make_desc(a_start, a_lbo, a_sbo);
        const uint64_t bdesc = make_desc(b_start, b_lbo, b_sbo);
        const uint32_t acc = ks > 0 ? 1u : 0u;
        if (!ts) {
          asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
                       " tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n"
                       ::"r"(tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
        } else {
          const uint32_t a_tmem = tmem + 64 + ks * 8;
          asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
                       " tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n}\n"
                       ::"r"(tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
        }
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    const bool ok = mbar_wait(smem_u32(&bar), 0);
    if (!ok && tid == 0) *status = 1;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (ok) {
#pragma unroll
      for (int c = 0; c < N / 16; ++c) {
        uint32_t r[16];
        const uint32_t addr = tmem + ((uint32_t)(warp * 32) << 16) + c * 16;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                       "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                     : "r"(addr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int i = 0; i < 16; ++i) D[(warp * 32 + lane) * N + c * 16 + i] = __uint_as_float(r[i]);
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128));
}

int main(int argc, char** argv) {
  const int variant = argc > 1 ? atoi(argv[1]) : 0;
  std::vector<float> A(M * KT), B(N * KT), D(M * N, -12345.f), want(M * N);
  for (int m = 0; m < M; ++m) for (int k = 0; k < KT; ++k) A[m * KT + k] = (float)((m * 7 + k * 13) % 17 - 8) / 8.f;
  for (int n = 0; n < N; ++n) for (int k = 0; k < KT; ++k) B[n * KT + k] = (float)((n * 5 + k * 3) % 13 - 6) / 4.f;
  for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) {
    double s = 0; for (int k = 0; k < KT; ++k) s += (double)A[m * KT + k] * B[n * KT + k];
    want[m * N + n] = (float)s;
  }
  float *dA, *dB, *dD; int* dS;
  CK(cudaMalloc(&dA, A.size() * 4)); CK(cudaMalloc(&dB, B.size() * 4)); CK(cudaMalloc(&dD, D.size() * 4)); CK(cudaMalloc(&dS, 4));
  CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dD, D.data(), D.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemset(dS, 0, 4));
  const size_t smem = (M * KT + N * KT) * 4 + 1024;
  probe<<<1, 128, smem>>>(dA, dB, dD, variant, dS);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  int st = 0;
  CK(cudaMemcpy(&st, dS, 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
  double maxerr = 0; int bad = -1, nbad = 0;
  if (variant == 7) {
    for (int m = 0; m < M; ++m) for (int k = 0; k < KT; ++k) {
      double e = fabs((double)D[m * N + k] - A[m * KT + k]);
      if (e > 1e-6) { if (bad < 0) bad = m * N + k; ++nbad; }
      if (e > maxerr) maxerr = e;
    }
  } else {
    for (int i = 0; i < M * N; ++i) {
      double e = fabs((double)D[i] - want[i]);
      if (e > 1e-4) { if (bad < 0) bad = i; ++nbad; }
      if (e > maxerr) maxerr = e;
    }
  }
  printf("variant %d: status %d max_err %.6g mismatches %d", variant, st, maxerr, nbad);
  if (bad >= 0) printf("  first bad (m=%d,n=%d) got %g want %g", bad / N, bad % N, D[bad], variant == 7 ? A[(bad / N) * KT + bad % N] : want[bad]);
  printf("  => %s\n", (st == 0 && nbad == 0) ? "PASS" : "FAIL");
  return 0;
}
