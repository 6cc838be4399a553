// ADLM_PREC_FP32 path of the prototype head: CUDA-core fp32 FMA kernels.
//
// These kernels evaluate the reference's formula in the reference's arithmetic class
//   d = relu(sum z^2 + (-2 z.p + sum p^2))      model.py:207-219
//   s = log((d+1)/(d+eps)) | -d                 model.py:231-237
//   logits = s W^T                              model.py:161-162
// and its backward, with everything between z and the outputs kept on chip (the reference makes
// ~10 HBM passes over the M x N map).  They are the exact-arithmetic mode of the
// library (any K <= 256, any N, C <= 128) and the on-device cross-check for the tcgen05 kernels
// in tc_head.cu, which are the fast path.
//
// Tiling: one CTA = 64 pixels of one image x all prototypes, walked in chunks of 64 prototypes.
// 256 threads; in the contraction thread (tx,ty) owns pixels {tx+16i} x prototypes {4ty+j}.
#include "common.cuh"

namespace adlm {

namespace {

constexpr int TM = 64;       // pixels per CTA tile
constexpr int TN = 64;       // prototypes per chunk
constexpr int NT = 256;      // threads per CTA
constexpr int PS = TN + 4;   // row stride of the transposed prototype chunk (keeps float4 alignment)
constexpr int SS = TN + 1;   // row stride of the similarity / gradient tiles (conflict-free columns)
constexpr int JMAX = ADLM_MAX_PROTOS_PER_CLASS;

__device__ __forceinline__ void load_z_tile(float* zs, const float* zb, int K, int Kq, int HW, int p0) {
  for (int idx = threadIdx.x; idx < Kq * TM; idx += NT) {
    const int k = idx >> 6, px = idx & 63, p = p0 + px;
    zs[idx] = (k < K && p < HW) ? __ldg(zb + (size_t)k * HW + p) : 0.f;
  }
}

// P[n0..n0+63][0..K) -> ps[k][n] (transposed).  Lane pattern (8 k x 4 n per warp) reads whole 32-byte
// sectors and writes 32 distinct banks.
__device__ __forceinline__ void load_p_chunk(float* ps, const float* P, int K, int Kq, int N, int n0) {
  const int kblocks = Kq >> 3;
  for (int idx = threadIdx.x; idx < TN * Kq; idx += NT) {
    const int kk = idx & 7, nn = (idx >> 3) & 3, rest = idx >> 5;
    const int kb = rest % kblocks, nb = rest / kblocks;
    const int n = nb * 4 + nn, k = kb * 8 + kk;
    ps[k * PS + n] = (n0 + n < N && k < K) ? __ldg(P + (size_t)(n0 + n) * K + k) : 0.f;
  }
}

__device__ __forceinline__ void load_w_chunk(float* ws, const float* W, int C, int rows, int N, int n0) {
  for (int idx = threadIdx.x; idx < rows * TN; idx += NT) {
    const int c = idx >> 6, n = idx & 63;
    ws[idx] = (c < C && n0 + n < N) ? __ldg(W + (size_t)c * N + n0 + n) : 0.f;
  }
}

__device__ __forceinline__ void tile_gemm(const float* zs, const float* ps, int K, int tx, int ty,
                                          float (&acc)[4][4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 4
  for (int k = 0; k < K; ++k) {
    float a[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = zs[k * TM + tx + 16 * i];
    const float4 b4 = *reinterpret_cast<const float4*>(ps + k * PS + ty * 4);
    const float b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
  }
}

// ------------------------------------------------------------------------------------ forward
template <int CPT>
__global__ void __launch_bounds__(NT)
k_simt_fwd(HeadShape s, const float* __restrict__ z, const float* __restrict__ P,
           const float* __restrict__ W, float* __restrict__ logits, float* __restrict__ dist,
           float* __restrict__ sim, int32_t* __restrict__ argmax, StepExtras ex, int tiles_per_img) {
  extern __shared__ __align__(16) float smem[];
  const int K = s.K, Kq = (K + 63) & ~63, N = s.N, C = s.C, HW = s.HW;
  float* zs = smem;
  float* ps = zs + Kq * TM;
  float* ss = ps + Kq * PS;
  float* ws = ss + TM * SS;             // [4*CPT][TN]
  float* ls = ws + 4 * CPT * TN;        // [TM][C+1]
  float* x2s = ls + TM * (C + 1);
  float* p2s = x2s + TM;

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int b = blockIdx.x / tiles_per_img, p0 = (blockIdx.x % tiles_per_img) * TM;
  const float* zb = z + (size_t)b * K * HW;

  load_z_tile(zs, zb, K, Kq, HW, p0);
  __syncthreads();
  if (tid < TM) {
    float a = 0.f;
    for (int k = 0; k < K; ++k) { const float v = zs[k * TM + tid]; a = fmaf(v, v, a); }
    x2s[tid] = a;
  }

  float lacc[CPT];
#pragma unroll
  for (int q = 0; q < CPT; ++q) lacc[q] = 0.f;
  const int lpx = tid & 63, lcg = tid >> 6;

  int lab[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = p0 + tx + 16 * i;
    lab[i] = (ex.own_dist != nullptr && p < HW) ? ex.target[(size_t)b * HW + p] : -1;
  }

  for (int n0 = 0; n0 < N; n0 += TN) {
    __syncthreads();
    load_p_chunk(ps, P, K, Kq, N, n0);
    load_w_chunk(ws, W, C, 4 * CPT, N, n0);
    __syncthreads();
    if (tid >= 64 && tid < 128) {
      const int n = tid - 64;
      float a = 0.f;
      for (int k = 0; k < K; ++k) { const float v = ps[k * PS + n]; a = fmaf(v, v, a); }
      p2s[n] = a;
    }
    float acc[4][4];
    tile_gemm(zs, ps, K, tx, ty, acc);
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int px = tx + 16 * i, p = p0 + px;
      const size_t m = (size_t)b * HW + p;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int nl = ty * 4 + j, n = n0 + nl;
        const float inter = fmaf(-2.f, acc[i][j], p2s[nl]);
        const float d = fmaxf(x2s[px] + inter, 0.f);
        float sv = act_fwd_precise(d, s.eps, s.act_mode);
        if (n >= N) sv = 0.f;
        ss[px * SS + nl] = sv;
        if (p < HW && n < N) {
          if (dist) dist[((size_t)b * N + n) * HW + p] = d;
          if (sim) sim[m * N + n] = sv;
          if (lab[i] >= 0 && lab[i] == __ldg(ex.proto_class + n)) {
            const int rk = __ldg(ex.proto_rank + n);      // < 0: beyond ADLM_MAX_PROTOS_PER_CLASS (flagged by the prologue)
            if (rk >= 0) ex.own_dist[m * JMAX + rk] = d;
          }
        }
      }
    }
    __syncthreads();
#pragma unroll 4
    for (int n = 0; n < TN; ++n) {
      const float sv = ss[lpx * SS + n];
#pragma unroll
      for (int q = 0; q < CPT; ++q) lacc[q] = fmaf(sv, ws[(lcg + 4 * q) * TN + n], lacc[q]);
    }
  }
  __syncthreads();
#pragma unroll
  for (int q = 0; q < CPT; ++q) {
    const int c = lcg + 4 * q;
    if (c < C) ls[lpx * (C + 1) + c] = lacc[q];
  }
  __syncthreads();
  float* lout = logits + ((size_t)b * HW + p0) * C;
  for (int idx = tid; idx < TM * C; idx += NT) {
    const int px = idx / C, c = idx - px * C;
    if (p0 + px < HW) lout[idx] = ls[px * (C + 1) + c];
  }
  if (argmax != nullptr && tid < TM && p0 + tid < HW) {
    int best = 0;
    float bv = ls[tid * (C + 1)];
    for (int c = 1; c < C; ++c) {
      const float v = ls[tid * (C + 1) + c];
      if (v > bv) { bv = v; best = c; }
    }
    argmax[(size_t)b * HW + p0 + tid] = best;
  }
}

// ------------------------------------------------------------------------------------ backward
// G[m][n] = ((dlogits W)[m][n] + dsim[m][n]) * ds/dd + ddist[m][n] + kld_grad,  zeroed where relu clipped
// dZ = 2 (rowsum(G) z - G P)     dP = 2 (colsum(G) P - G^T Z)     dW = dlogits^T S
template <int KB>
__global__ void __launch_bounds__(NT)
k_simt_bwd(HeadShape s, const float* __restrict__ z, const float* __restrict__ P,
           const float* __restrict__ W, const float* __restrict__ dlogits,
           const float* __restrict__ ddist, const float* __restrict__ dsim,
           float* __restrict__ dz, float* __restrict__ dP, float* __restrict__ dW,
           StepExtras ex, int tiles_per_img) {
  extern __shared__ __align__(16) float smem[];
  const int K = s.K, N = s.N, C = s.C, HW = s.HW;
  constexpr int Kq = KB * 64;
  float* zs = smem;                     // [Kq][TM]
  float* ps = zs + Kq * TM;             // [Kq][PS]
  float* gs = ps + Kq * PS;             // [TM][SS]   G tile
  float* ss = gs + TM * SS;             // [TM][SS]   S tile
  float* ws = ss + TM * SS;             // [C][TN]
  float* dls = ws + C * TN;             // [TM][C+1]
  float* x2s = dls + TM * (C + 1);
  float* p2s = x2s + TM;

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int b = blockIdx.x / tiles_per_img, p0 = (blockIdx.x % tiles_per_img) * TM;
  const float* zb = z + (size_t)b * K * HW;

  load_z_tile(zs, zb, K, Kq, HW, p0);
  if (dlogits != nullptr) {
    const float* dl = dlogits + ((size_t)b * HW + p0) * C;
    for (int idx = tid; idx < TM * C; idx += NT) {
      const int px = idx / C, c = idx - px * C;
      dls[px * (C + 1) + c] = (p0 + px < HW) ? __ldg(dl + idx) : 0.f;
    }
  }
  __syncthreads();
  if (tid < TM) {
    float a = 0.f;
    for (int k = 0; k < K; ++k) { const float v = zs[k * TM + tid]; a = fmaf(v, v, a); }
    x2s[tid] = a;
  }

  int lab[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = p0 + tx + 16 * i;
    lab[i] = (ex.own_grad != nullptr && p < HW) ? ex.target[(size_t)b * HW + p] : -1;
  }

  float dzacc[KB][4][4];
  float rsum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int kb = 0; kb < KB; ++kb)
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) dzacc[kb][i][j] = 0.f;

  for (int n0 = 0; n0 < N; n0 += TN) {
    __syncthreads();
    load_p_chunk(ps, P, K, Kq, N, n0);
    if (dlogits != nullptr) load_w_chunk(ws, W, C, C, N, n0);
    __syncthreads();
    if (tid >= 64 && tid < 128) {
      const int n = tid - 64;
      float a = 0.f;
      for (int k = 0; k < K; ++k) { const float v = ps[k * PS + n]; a = fmaf(v, v, a); }
      p2s[n] = a;
    }
    float acc[4][4];
    tile_gemm(zs, ps, K, tx, ty, acc);
    float g[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) g[i][j] = 0.f;
    if (dlogits != nullptr) {
      for (int c = 0; c < C; ++c) {
        float a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = dls[(tx + 16 * i) * (C + 1) + c];
        const float4 w4 = *reinterpret_cast<const float4*>(ws + c * TN + ty * 4);
        const float wv[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) g[i][j] = fmaf(a[i], wv[j], g[i][j]);
      }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int px = tx + 16 * i, p = p0 + px;
      const size_t m = (size_t)b * HW + p;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int nl = ty * 4 + j, n = n0 + nl;
        float G = 0.f, sv = 0.f;
        if (p < HW && n < N) {
          const float pre = x2s[px] + fmaf(-2.f, acc[i][j], p2s[nl]);
          const float d = fmaxf(pre, 0.f);
          sv = act_fwd_precise(d, s.eps, s.act_mode);
          float up = g[i][j];
          if (dsim) up += __ldg(dsim + m * N + n);
          G = up * act_grad(d, s.eps, s.act_mode);
          if (ddist) G += __ldg(ddist + ((size_t)b * N + n) * HW + p);
          if (lab[i] >= 0 && lab[i] == __ldg(ex.proto_class + n)) {
            const int rk = __ldg(ex.proto_rank + n);
            if (rk >= 0) G += __ldg(ex.own_grad + m * JMAX + rk);
          }
          if (!(pre > 0.f)) G = 0.f;
        }
        gs[px * SS + nl] = G;
        ss[px * SS + nl] = sv;
      }
    }
    __syncthreads();

    if (dz != nullptr) {
#pragma unroll 2
      for (int n = 0; n < TN; ++n) {
        float gv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { gv[i] = gs[(tx + 16 * i) * SS + n]; rsum[i] += gv[i]; }
#pragma unroll
        for (int kb = 0; kb < KB; ++kb) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float pv = ps[(kb * 64 + ty * 4 + j) * PS + n];
#pragma unroll
            for (int i = 0; i < 4; ++i) dzacc[kb][i][j] = fmaf(gv[i], pv, dzacc[kb][i][j]);
          }
        }
      }
    }
    if (dP != nullptr) {
      // thread (tn, tk): prototypes {tn+16i}, depth {64kb + 4tk + j}
      const int tn = tx, tk = ty;
      float csum[4] = {0.f, 0.f, 0.f, 0.f};
      for (int px = 0; px < TM; ++px) {
#pragma unroll
        for (int i = 0; i < 4; ++i) csum[i] += gs[px * SS + tn + 16 * i];
      }
#pragma unroll 1
      for (int kb = 0; kb < KB; ++kb) {
        float dp[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) dp[i][j] = 0.f;
#pragma unroll 2
        for (int px = 0; px < TM; ++px) {
          float gv[4], zv[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) gv[i] = gs[px * SS + tn + 16 * i];
#pragma unroll
          for (int j = 0; j < 4; ++j) zv[j] = zs[(kb * 64 + tk * 4 + j) * TM + px];
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) dp[i][j] = fmaf(gv[i], zv[j], dp[i][j]);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int n = n0 + tn + 16 * i;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int k = kb * 64 + tk * 4 + j;
            if (n < N && k < K) {
              const float v = 2.f * (csum[i] * ps[k * PS + tn + 16 * i] - dp[i][j]);
              atomicAdd(dP + (size_t)n * K + k, v);
            }
          }
        }
      }
    }
    if (dW != nullptr && dlogits != nullptr) {
      for (int idx = tid; idx < C * TN; idx += NT) {
        const int c = idx >> 6, nl = idx & 63;
        if (n0 + nl < N) {
          float a = 0.f;
#pragma unroll 4
          for (int px = 0; px < TM; ++px) a = fmaf(dls[px * (C + 1) + c], ss[px * SS + nl], a);
          atomicAdd(dW + (size_t)c * N + n0 + nl, a);
        }
      }
    }
  }

  if (dz != nullptr) {
    float* dzb = dz + (size_t)b * K * HW;
#pragma unroll
    for (int kb = 0; kb < KB; ++kb)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = kb * 64 + ty * 4 + j;
        if (k < K) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int px = tx + 16 * i, p = p0 + px;
            if (p < HW) dzb[(size_t)k * HW + p] = 2.f * (rsum[i] * zs[k * TM + px] - dzacc[kb][i][j]);
          }
        }
      }
  }
}

template <typename Kern>
int set_smem(Kern kern, size_t bytes) {
  if (bytes > 227 * 1024) {
    set_error("simt head: %zu bytes of shared memory needed (> 227 KB)", bytes);
    return ADLM_E_UNSUPPORTED;
  }
  ADLM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return ADLM_OK;
}

}  // namespace

int simt_head_fwd(const HeadShape& s, const float* z, const float* P, const float* W,
                  float* logits, float* dist, float* sim, int32_t* argmax,
                  const StepExtras& ex, cudaStream_t st) {
  const int Kq = (s.K + 63) & ~63;
  const int tiles_per_img = cdiv(s.HW, TM);
  const long long grid = (long long)s.B * tiles_per_img;
  ADLM_CHECK_SUPPORTED(grid < (1ll << 31), "too many pixel tiles");
  const int cpt = (s.C <= 32) ? 8 : 32;
  const size_t bytes = sizeof(float) * ((size_t)Kq * TM + (size_t)Kq * PS + TM * SS + 4 * cpt * TN +
                                        (size_t)TM * (s.C + 1) + TM + TN);
  int rc;
  KernelTimer timer(ADLM_K_HEAD_FWD, st);
  if (cpt == 8) {
    if ((rc = set_smem(k_simt_fwd<8>, bytes)) != ADLM_OK) return rc;
    k_simt_fwd<8><<<(unsigned)grid, NT, bytes, st>>>(s, z, P, W, logits, dist, sim, argmax, ex, tiles_per_img);
  } else {
    if ((rc = set_smem(k_simt_fwd<32>, bytes)) != ADLM_OK) return rc;
    k_simt_fwd<32><<<(unsigned)grid, NT, bytes, st>>>(s, z, P, W, logits, dist, sim, argmax, ex, tiles_per_img);
  }
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

int simt_head_bwd(const HeadShape& s, const float* z, const float* P, const float* W,
                  const float* dlogits, const float* ddist, const float* dsim,
                  float* dz, float* dP, float* dW, const StepExtras& ex, cudaStream_t st) {
  const int KB = cdiv(s.K, 64);
  const int Kq = KB * 64;
  const int tiles_per_img = cdiv(s.HW, TM);
  const long long grid = (long long)s.B * tiles_per_img;
  ADLM_CHECK_SUPPORTED(grid < (1ll << 31), "too many pixel tiles");
  const size_t bytes = sizeof(float) * ((size_t)Kq * TM + (size_t)Kq * PS + 2 * TM * SS + (size_t)s.C * TN +
                                        (size_t)TM * (s.C + 1) + TM + TN);
  if (dP) ADLM_CUDA(cudaMemsetAsync(dP, 0, sizeof(float) * (size_t)s.N * s.K, st));
  if (dW) ADLM_CUDA(cudaMemsetAsync(dW, 0, sizeof(float) * (size_t)s.C * s.N, st));
  int rc;
  KernelTimer timer(ADLM_K_HEAD_BWD, st);
#define ADLM_BWD_CASE(KBV)                                                                         \
  case KBV:                                                                                        \
    if ((rc = set_smem(k_simt_bwd<KBV>, bytes)) != ADLM_OK) return rc;                             \
    k_simt_bwd<KBV><<<(unsigned)grid, NT, bytes, st>>>(s, z, P, W, dlogits, ddist, dsim, dz, dP, dW, ex, \
                                                       tiles_per_img);                             \
    break;
  switch (KB) {
    ADLM_BWD_CASE(1)
    ADLM_BWD_CASE(2)
    ADLM_BWD_CASE(3)
    ADLM_BWD_CASE(4)
    default:
      set_error("simt_head_bwd: K=%d > %d", s.K, ADLM_MAX_K);
      return ADLM_E_UNSUPPORTED;
  }
#undef ADLM_BWD_CASE
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

}  // namespace adlm
