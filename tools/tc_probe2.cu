// Bring-up probe 2 (not part of the library): 128-byte-swizzled operand layouts and the M=64 accumulator layout.
//   tc_probe2 <variant>
//   10: A and B K-major SWIZZLE_128B            D[128x64] = A[128x64] B[64x64]^T
//   11: A MN-major SWIZZLE_128B, B K-major SW128 D[128x64] = A[128x16] B[64x16]^T   (A stored [k][m], m contiguous)
//   12: A K-major SW128, B MN-major SWIZZLE_128B                                    (B stored [k][n], n contiguous)
//   13: M=64 (A,B K-major no swizzle): prints which TMEM lane holds each accumulator row
//   14: A MN-major SW128 with M=64
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 2; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;   // 2 = SWIZZLE_128B
  return d;
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
// row-major [rows][cols] fp32 tile with 128-byte swizzle: cols in atoms of 32 floats; atom (ca, rg) at
// ca*rows*128 + rg*1024 bytes; inside: row r%8 at +128*(r%8), 16-byte chunk (c%32)/4 XOR (r%8)
__device__ __forceinline__ int off_sw128(int r, int c, int rows) {
  const int ca = c >> 5, ci = c & 31, chunk = (ci >> 2) ^ (r & 7);
  return ca * rows * 32 + (r >> 3) * 256 + (r & 7) * 32 + chunk * 4 + (ci & 3);
}
__device__ __forceinline__ int off_kmajor_ns(int r, int k, int R) { return (k >> 2) * R * 4 + r * 4 + (k & 3); }

struct Cfg { int M, N, KT; };

__global__ void __launch_bounds__(128) probe(const float* __restrict__ A, const float* __restrict__ B,
                                             float* __restrict__ D, int variant, Cfg cfg, int* __restrict__ status) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int M = cfg.M, N = cfg.N, KT = cfg.KT;
  float* sA = reinterpret_cast<float*>(smem);
  float* sB = sA + 128 * 64;    // 32 KB reserved for A
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool a_mn = (variant == 11 || variant == 14), b_mn = (variant == 12), ns = (variant == 13);

  for (int idx = tid; idx < M * KT; idx += 128) {
    const int m = idx / KT, k = idx % KT;
    int off;
    if (ns) off = off_kmajor_ns(m, k, M);
    else if (a_mn) off = off_sw128(k, m, KT);      // rows = k, cols = m
    else off = off_sw128(m, k, M);                 // rows = m, cols = k
    sA[off] = A[idx];
  }
  for (int idx = tid; idx < N * KT; idx += 128) {
    const int n = idx / KT, k = idx % KT;
    int off;
    if (ns) off = off_kmajor_ns(n, k, N);
    else if (b_mn) off = off_sw128(k, n, KT);
    else off = off_sw128(n, k, N);
    sB[off] = B[idx];
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(64));
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
  // pre-fill the accumulator with a sentinel so untouched lanes are visible (variant 13/14)
  {
    uint32_t v[16];
    for (int i = 0; i < 16; ++i) v[i] = __float_as_uint(-777.f);
    for (int c = 0; c < 4; ++c) {
      const uint32_t addr = tmem + ((uint32_t)(warp * 32) << 16) + c * 16;
      asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                   ::"r"(addr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                     "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  if (tid == 0) {
    const uint32_t idesc = make_idesc(M, N, a_mn ? 1 : 0, b_mn ? 1 : 0);
    for (int ks = 0; ks < KT / 8; ++ks) {
      uint64_t adesc, bdesc;
      if (ns) {
        adesc = make_desc(smem_u32(sA) + ks * 2 * M * 16, M * 16, 128, 0);
        bdesc = make_desc(smem_u32(sB) + ks * 2 * N * 16, N * 16, 128, 0);
      } else {
        // K-major SW128: k-atom (32 floats) stride = rows*128 B, 8-row group stride (SBO) = 1024 B,
        //   k-step of 8 floats inside the atom = +32 B on the start address
        // MN-major SW128: rows = k; one k-step = one 8-row group (+1024 B = SBO); MN atoms (32 floats) at LBO = KT*128 B
        if (!a_mn) adesc = make_desc(smem_u32(sA) + (ks >> 2) * M * 128 + (ks & 3) * 32, 16, 1024, 2);
        else       adesc = make_desc(smem_u32(sA) + ks * 1024, KT * 128, 1024, 2);
        if (!b_mn) bdesc = make_desc(smem_u32(sB) + (ks >> 2) * N * 128 + (ks & 3) * 32, 16, 1024, 2);
        else       bdesc = make_desc(smem_u32(sB) + ks * 1024, KT * 128, 1024, 2);
      }
      const uint32_t acc = ks > 0 ? 1u : 0u;
      asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
                   " tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n"
                   ::"r"(tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
  }
  const bool ok = mbar_wait(smem_u32(&bar), 0);
  if (!ok && tid == 0) *status = 1;
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (ok) {
    for (int c = 0; c < 4; ++c) {
      uint32_t r[16];
      const uint32_t addr = tmem + ((uint32_t)(warp * 32) << 16) + c * 16;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                   : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                     "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                   : "r"(addr));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      for (int i = 0; i < 16; ++i) D[(warp * 32 + lane) * 64 + c * 16 + i] = __uint_as_float(r[i]);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64));
}

int main(int argc, char** argv) {
  const int variant = argc > 1 ? atoi(argv[1]) : 10;
  Cfg cfg{128, 64, 64};
  if (variant == 11 || variant == 12) cfg.KT = 16;
  if (variant == 13) { cfg.M = 64; cfg.KT = 16; }
  if (variant == 14) { cfg.M = 64; cfg.KT = 16; }
  const int M = cfg.M, N = cfg.N, KT = cfg.KT;
  std::vector<float> A(M * KT), B(N * KT), D(128 * 64, -12345.f), want(M * N);
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
  const size_t smem = 128 * 64 * 4 + 64 * 64 * 4 + 1024;
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  probe<<<1, 128, smem>>>(dA, dB, dD, variant, cfg, dS);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  int st = 0;
  CK(cudaMemcpy(&st, dS, 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
  if (M == 64) {
    // find, for every accumulator row, the lane that holds it
    printf("variant %d: status %d; row -> lane: ", variant, st);
    int found = 0;
    for (int m = 0; m < M; ++m) {
      int where = -1;
      for (int l = 0; l < 128 && where < 0; ++l) {
        bool eq = true;
        for (int n = 0; n < N; ++n) if (fabs(D[l * 64 + n] - want[m * N + n]) > 1e-4) { eq = false; break; }
        if (eq) where = l;
      }
      if (where >= 0) ++found;
      if (m < 8 || m % 16 == 0 || m % 16 == 15) printf("%d->%d ", m, where);
    }
    int untouched = 0;
    for (int l = 0; l < 128; ++l) if (D[l * 64] == -777.f) ++untouched;
    printf(" | rows found %d/64, untouched lanes %d  => %s\n", found, untouched, found == 64 ? "PASS" : "FAIL");
    return 0;
  }
  double maxerr = 0; int bad = -1, nbad = 0;
  for (int i = 0; i < M * N; ++i) {
    double e = fabs((double)D[i] - want[i]);
    if (e > 1e-4) { if (bad < 0) bad = i; ++nbad; }
    if (e > maxerr) maxerr = e;
  }
  printf("variant %d: status %d max_err %.6g mismatches %d", variant, st, maxerr, nbad);
  if (bad >= 0) printf("  first bad (m=%d,n=%d) got %g want %g", bad / N, bad % N, D[bad], want[bad]);
  printf("  => %s\n", (st == 0 && nbad == 0) ? "PASS" : "FAIL");
  return 0;
}
