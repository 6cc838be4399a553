// ADLM_PREC_TF32 / ADLM_PREC_TF32X3 path: the prototype head on the 5th-generation tensor cores.
//
// Forward (k_tc_fwd), per 128-pixel tile, one thread per pixel (= one TMEM lane):
//   convert   z[k][px] (NCHW, coalesced LDG, prefetched one tile ahead) -> centred (z - 1/2; squared L2
//             distance is translation invariant), rounded to nearest tf32 (hi [+ lo for 3xTF32]) ->
//             shared-memory A operand [K/4][128 px][4 k]; |z|^2 accumulated in a register.
//             (TMA cannot fetch this tile: the NCHW row pitch h*w*4 B is not a multiple of 16 for the
//             reference's odd maps — 65x65, 129x257 — and a register pass is needed anyway to round/split.)
//   MMA1      D1[128 x Npad] = A . Pk^T     tcgen05.mma kind::tf32, accumulator in TMEM
//   epilogue1 tcgen05.ld -> d = max(|z|^2 - 2 acc + |p|^2, 0) -> s = log((d+1)/(d+eps)); optional streaming
//             stores of d (NCHW), s, own-class distances; s is written back over D1 (tcgen05.st)
//   MMA2      D2[128 x Cpad] = S . W^T      A operand straight from TMEM (.ts form): the similarity matrix
//             never touches shared or global memory
//   epilogue2 logits (B,h,w,C) + per-pixel argmax
// Two warpgroups per CTA run this pipeline on alternating tiles, so one group's MUFU-bound epilogue
// overlaps the other's loads and MMAs; the grid is persistent (one CTA per SM).
// 3xTF32 keeps fp32-class distances (hi*lo + lo*hi + hi*hi) and does the last layer in fp32 FMA.
#include <cstdlib>
#include "tc_head.cuh"
#include "tc_common.cuh"

namespace adlm {

namespace {

using namespace tc;

constexpr int TILE = 128;
constexpr int JMAX = ADLM_MAX_PROTOS_PER_CLASS;
constexpr float LN2 = 0.6931471805599453f;
constexpr float CENTRE = 0.5f;

// -DADLM_PHASE_CLOCKS: CTA 0 records clock64() at the phase boundaries of its first tiles (tools/phase_clocks.py)
#ifdef ADLM_PHASE_CLOCKS
__device__ long long g_phase[2][64];
#define ADLM_PHASE(i)                                                        \
  do {                                                                       \
    if (blockIdx.x == 0 && wt == 0 && (i) < 64) g_phase[wg][(i)] = clock64(); \
  } while (0)
#else
#define ADLM_PHASE(i)
#endif

template <int KP, bool X3>
struct FwdSmem {
  static constexpr int NWG = X3 ? 1 : 2;
  static constexpr int PARTS = X3 ? 2 : 1;
  // byte offsets
  static __host__ __device__ size_t pk(int part, int Npad) { return (size_t)part * KP * Npad * 4; }
  static __host__ __device__ size_t wk(int Npad) { return (size_t)PARTS * KP * Npad * 4; }
  // 3xTF32: [Npad][32] fp32 for the CUDA-core last layer, or hi + lo K-major operands (2 x Npad x Cpad) for the
  // tensor-core one — the larger of the two, so one layout serves both (Cpad <= 32)
  static __host__ __device__ size_t wk_bytes(int Npad, int Cpad, bool tcl = false) {
    return (size_t)Npad * (X3 ? (tcl ? 2 * Cpad : 32) : Cpad) * 4;
  }
  static __host__ __device__ size_t abuf(int wg, int part, int Npad, int Cpad, bool tcl = false) {
    return wk(Npad) + wk_bytes(Npad, Cpad, tcl) + ((size_t)wg * PARTS + part) * KP * TILE * 4;
  }
  static __host__ __device__ size_t misc(int Npad, int Cpad, bool tcl = false) { return abuf(NWG, 0, Npad, Cpad, tcl); }
  static __host__ __device__ size_t total(int Npad, int Cpad, bool tcl = false) {
    // p2 [Npad] f32, class [Npad] i32, rank [Npad] i32, chunk class masks [Npad/16] u32, barriers, tmem base
    return misc(Npad, Cpad, tcl) + (size_t)Npad * 12 + (size_t)(Npad / 16) * 4 + 64 + 16;
  }
};

template <int KP, bool X3>
__global__ void __launch_bounds__(FwdSmem<KP, X3>::NWG * 128, 1) k_tc_fwd(const TcFwdParams p) {
  using L = FwdSmem<KP, X3>;
  constexpr int NWG = L::NWG;
  constexpr int NTHREADS = NWG * 128;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int Npad = p.Npad, Cpad = p.Cpad, K = p.K, N = p.N, C = p.C, HW = p.HW;
  float* Pk_hi = reinterpret_cast<float*>(smem + L::pk(0, Npad));
  float* Pk_lo = reinterpret_cast<float*>(smem + L::pk(X3 ? 1 : 0, Npad));
  float* Wk = reinterpret_cast<float*>(smem + L::wk(Npad));
  const bool tcl = X3 && p.x3_tensor_last != 0;          // 3xTF32 with the last layer on the tensor cores
  float* Wk_lo = Wk + (size_t)Npad * Cpad;               // (tcl) low parts, same K-major layout
  uint8_t* misc = smem + L::misc(Npad, Cpad, tcl);
  float* p2s = reinterpret_cast<float*>(misc);
  int* pcs = reinterpret_cast<int*>(misc + (size_t)Npad * 4);
  int* prs = reinterpret_cast<int*>(misc + (size_t)Npad * 8);
  uint32_t* cmask = reinterpret_cast<uint32_t*>(misc + (size_t)Npad * 12);
  uint64_t* bars = reinterpret_cast<uint64_t*>(misc + (size_t)Npad * 12 + (size_t)(Npad / 16) * 4 + ((16 - ((Npad / 16) * 4) % 16) % 16));
  uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(bars + 4);

  const int tid = threadIdx.x, wg = tid >> 7, wt = tid & 127, warp_in_wg = wt >> 5;
  const bool want_own = p.own_dist != nullptr;

  ADLM_PHASE(0);
  // ---- one-time setup: TMEM, barriers, prototype / last-layer operands ----
  if (tid < 32) tmem_alloc(smem_u32(tmem_base_s), X3 ? 512 : NWG * 256);
  if (tid == 32) {
    for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bars[i]), 1);
    mbar_init_fence();
  }
  // the first tile's features start travelling before the operand staging below
  const int stride = (int)gridDim.x * NWG;
  int tile = (int)blockIdx.x + wg * (int)gridDim.x;
  const int Mi = (int)p.M;
  float zr[KP];
  auto prefetch = [&](int t) {
    const int m = t * TILE + wt;
    if (t < p.num_tiles && m < Mi) {
      const int b = m / HW;
      const float* src = p.z + ((size_t)b * K) * HW + (m - b * HW);
      if (K == KP) {
#pragma unroll
        for (int k = 0; k < KP; ++k) { zr[k] = __ldg(src); src += HW; }
      } else {
#pragma unroll
        for (int k = 0; k < KP; ++k) zr[k] = (k < K) ? __ldg(src + (size_t)k * HW) : CENTRE;
      }
    } else {
#pragma unroll
      for (int k = 0; k < KP; ++k) zr[k] = CENTRE;
    }
  };
  prefetch(tile);

  // Pk: [KP/4][Npad][4], centred and rounded (3xTF32: hi and lo = rna(x - hi))
  stage_rows_kmajor<KP, X3>(Pk_hi, Pk_lo, p.P, N, K, Npad, CENTRE, tid, NTHREADS);
  if (!X3) {
    // Wk: B operand of MMA2, rows = classes, K index = prototype: [Npad/4][Cpad][4]; ln 2 folded in (the
    // epilogue produces similarities in log2 units).  Padding first (disjoint from the valid entries), then
    // the valid entries streamed coalesced from W [C][N].
    for (int idx = tid; idx < Npad * Cpad; idx += NTHREADS) {
      const int n = idx / Cpad, c = idx - n * Cpad;
      if (!(c < C && n < N)) Wk[(size_t)(n >> 2) * Cpad * 4 + c * 4 + (n & 3)] = 0.f;
    }
    const float wscale = p.act_mode == ADLM_ACT_LOG ? LN2 : 1.f;
    const int CN = C * N;
    for (int i0 = 0; i0 < CN; i0 += NTHREADS * 4) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int idx = i0 + u * NTHREADS + tid;
        v[u] = idx < CN ? __ldg(p.W + idx) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int idx = i0 + u * NTHREADS + tid;
        if (idx < CN) {
          const int c = idx / N, n = idx - c * N;
          Wk[(size_t)(n >> 2) * Cpad * 4 + c * 4 + (n & 3)] = rna_tf32_mask(wscale * v[u]);
        }
      }
    }
  } else if (tcl) {
    // hi + lo K-major B operands of the last-layer MMAs: [Npad/4][Cpad][4] each, w = hi + lo with hi = rna(w),
    // lo = rna(w - hi); the similarities stay in natural-log units here (no ln 2 folded in)
    for (int idx = tid; idx < Npad * Cpad; idx += NTHREADS) {
      const int n = idx / Cpad, c = idx - n * Cpad;
      const float wv = (c < C && n < N) ? __ldg(p.W + (size_t)c * N + n) : 0.f;
      const float hi = rna_tf32_mask(wv);
      const size_t o = (size_t)(n >> 2) * Cpad * 4 + c * 4 + (n & 3);
      Wk[o] = hi;
      Wk_lo[o] = rna_tf32_mask(wv - hi);
    }
  } else {
    // plain fp32 [Npad][32] for the CUDA-core last layer
    for (int idx = tid; idx < Npad * 32; idx += NTHREADS) {
      const int n = idx >> 5, c = idx & 31;
      if (!(c < C && n < N)) Wk[idx] = 0.f;
    }
    const int CN = C * N;
    for (int idx = tid; idx < CN; idx += NTHREADS) {
      const int c = idx / N, n = idx - c * N;
      Wk[n * 32 + c] = __ldg(p.W + idx);
    }
  }
  __syncthreads();
  for (int n = tid; n < Npad; n += NTHREADS) {
    // 1xTF32: |p|^2 of the ROUNDED vector (so that z == p gives d == 0); 3xTF32: hi + lo is the fp32 value
    float a = 0.f;
    for (int k4 = 0; k4 < KP / 4; ++k4) {
      const float4 h4 = *reinterpret_cast<const float4*>(Pk_hi + ((size_t)k4 * Npad + n) * 4);
      float4 v = h4;
      if (X3) {
        const float4 l4 = *reinterpret_cast<const float4*>(Pk_lo + ((size_t)k4 * Npad + n) * 4);
        v.x += l4.x; v.y += l4.y; v.z += l4.z; v.w += l4.w;
      }
      a = fmaf(v.x, v.x, a); a = fmaf(v.y, v.y, a); a = fmaf(v.z, v.z, a); a = fmaf(v.w, v.w, a);
    }
    p2s[n] = a;
    // rank < 0: the prototype lies beyond ADLM_MAX_PROTOS_PER_CLASS (the prologue kernel flags the step as failed)
    const int rk = (want_own && n < N) ? p.proto_rank[n] : -1;
    pcs[n] = rk >= 0 ? p.proto_class[n] : -1;
    prs[n] = rk >= 0 ? rk : 0;
  }
  __syncthreads();
  for (int c = tid; c < Npad / 16; c += NTHREADS) {
    uint32_t m = 0u;
    for (int i = 0; i < 16; ++i) {
      const int pc = pcs[c * 16 + i];
      if (pc >= 0 && pc < 32) m |= 1u << pc;
    }
    cmask[c] = m;
  }
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();

  const uint32_t tmem_base = *tmem_base_s;
  const uint32_t d1 = tmem_base + wg * 256;              // [128 lanes x Npad cols]
  const uint32_t slo = d1 + Npad;                        // (tcl) low parts of the similarities [128 x Npad]
  const uint32_t d2 = d1 + (tcl ? 2 * Npad : Npad);      // [128 lanes x Cpad cols]
  const uint32_t lane_field = (uint32_t)(warp_in_wg * 32) << 16;
  float* Ahi = reinterpret_cast<float*>(smem + L::abuf(wg, 0, Npad, Cpad, tcl));
  float* Alo = reinterpret_cast<float*>(smem + L::abuf(wg, X3 ? 1 : 0, Npad, Cpad, tcl));
  const uint32_t bar_d1 = smem_u32(&bars[wg * 2]), bar_d2 = smem_u32(&bars[wg * 2 + 1]);
  const uint32_t idesc1 = make_idesc_tf32(TILE, Npad), idesc2 = make_idesc_tf32(TILE, Cpad);
  const int ksteps = (K + 7) >> 3;

  ADLM_PHASE(1);
  // Ping-pong of the two warpgroups: the MUFU-bound epilogue 1 is entered with a token (named barriers 3 + wg),
  // so one group's epilogue overlaps the other's loads, MMAs and stores instead of both fighting for the
  // MUFU pipe and then both idling.  n_mine / n_other: tiles of this / the other warpgroup.
  auto tiles_of = [&](int g) {
    const int first = (int)blockIdx.x + g * (int)gridDim.x;
    return first < p.num_tiles ? (p.num_tiles - 1 - first) / stride + 1 : 0;
  };
  const int n_mine = tiles_of(wg), n_other = NWG == 2 ? tiles_of(1 - wg) : 0;
  if (NWG == 2 && wg == 1 && n_other > 0) named_bar_arrive(3, 256);   // WG0 holds the token first
  int it = 0;
  (void)n_mine;
  uint32_t parity = 0;
  int ph = 2;
  (void)ph;
  for (; tile < p.num_tiles; tile += stride, parity ^= 1u, ++it) {
    ADLM_PHASE(ph); ++ph;                      // +0 tile start
    const int m = tile * TILE + wt;
    const bool valid = m < Mi;
    const int b = valid ? m / HW : 0;
    const int pix = valid ? m - b * HW : 0;

    // ---- convert: registers -> tf32 operand tile; |z|^2 ----
    float x2 = 0.f;
#pragma unroll
    for (int k4 = 0; k4 < KP / 4; ++k4) {
      float hi[4], lo[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float v = zr[k4 * 4 + j] - CENTRE;
        hi[j] = rna_tf32_mask(v);
        lo[j] = X3 ? rna_tf32_mask(v - hi[j]) : 0.f;
        const float u = X3 ? v : hi[j];
        x2 = fmaf(u, u, x2);
      }
      *reinterpret_cast<float4*>(Ahi + ((size_t)k4 * TILE + wt) * 4) = make_float4(hi[0], hi[1], hi[2], hi[3]);
      if (X3) *reinterpret_cast<float4*>(Alo + ((size_t)k4 * TILE + wt) * 4) = make_float4(lo[0], lo[1], lo[2], lo[3]);
    }
    fence_async_smem();
    fence_before();
    named_bar_sync(1 + wg, 128);
    ADLM_PHASE(ph); ++ph;                      // +1 converted, barrier passed
    if (wt == 0) {
      fence_after();
      const uint32_t a_hi = smem_u32(Ahi), a_lo = smem_u32(Alo), b_hi = smem_u32(Pk_hi), b_lo = smem_u32(Pk_lo);
      uint32_t acc = 0u;
      for (int ks = 0; ks < ksteps; ++ks) {
        const uint32_t ao = ks * 2 * TILE * 16, bo = ks * 2 * Npad * 16;
        if (X3) {
          mma_ss(d1, make_desc(a_hi + ao, TILE * 16, 128), make_desc(b_lo + bo, Npad * 16, 128), idesc1, acc);
          mma_ss(d1, make_desc(a_lo + ao, TILE * 16, 128), make_desc(b_hi + bo, Npad * 16, 128), idesc1, 1u);
          acc = 1u;
        }
        mma_ss(d1, make_desc(a_hi + ao, TILE * 16, 128), make_desc(b_hi + bo, Npad * 16, 128), idesc1, acc);
        acc = 1u;
      }
      mma_commit(bar_d1);
    }
    // next tile's features travel while the tensor core and the epilogue work
    ADLM_PHASE(ph); ++ph;                      // +2 MMA1 issued
    prefetch(tile + stride);
    ADLM_PHASE(ph); ++ph;                      // +3 prefetch issued

    int lab = -1;
    if (want_own && valid) lab = p.target[m];
    const bool lab_ok = lab >= 0 && lab < 32;

    mbar_wait(bar_d1, parity);
    fence_after();
    ADLM_PHASE(ph); ++ph;                      // +4 MMA1 done

    // ---- epilogue 1: distances -> similarities ----
    if (NWG == 2) named_bar_sync(3 + wg, 256);
    float lacc[X3 ? 32 : 1];
    if (X3) {
#pragma unroll
      for (int c = 0; c < 32; ++c) lacc[X3 ? c : 0] = 0.f;
    }
    const uint32_t t1 = d1 + lane_field;
    const bool act_log = p.act_mode == ADLM_ACT_LOG;
    const float eps = p.eps;
    const int nch = Npad / 16;
    auto chunk = [&](int ch, const uint32_t (&r)[16]) {
      float xp[16];
#pragma unroll
      for (int v4 = 0; v4 < 4; ++v4) {
        const float4 q = *reinterpret_cast<const float4*>(p2s + ch * 16 + v4 * 4);
        xp[v4 * 4] = x2 + q.x; xp[v4 * 4 + 1] = x2 + q.y; xp[v4 * 4 + 2] = x2 + q.z; xp[v4 * 4 + 3] = x2 + q.w;
      }
      float dv[16];
      uint32_t sv_bits[16];
      float sv[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float d = fmaxf(fmaf(-2.f, __uint_as_float(r[i]), xp[i]), 0.f);
        dv[i] = d;
        float s;
        // 1xTF32: similarity in log2 units (ln 2 is folded into the last-layer operand Wk)
#if defined(ADLM_EXP) && ADLM_EXP == 1
        if (act_log) s = (d + 1.f) - 0.5f * (d + eps);
#else
        // 3xTF32: log(1 + (1-eps)/(d+eps)) with the hardware reciprocal and log2 (each ~1 ulp; the form has no
        // cancellation, so the similarity is good to ~3e-7 absolute — logf + a division cost ~40 instructions here,
        // most of this kernel's time)
        if (act_log) s = X3 ? LN2 * lg2_approx(fmaf(1.f - eps, rcp_approx(d + eps), 1.f)) : lg2_approx(d + 1.f) - lg2_approx(d + eps);
#endif
        else s = -d;
        sv[i] = s;
        sv_bits[i] = __float_as_uint(rna_tf32_fast(s));
      }
      if (!X3) {
        tmem_st16(t1 + ch * 16, sv_bits);
      } else if (tcl) {
        // s = hi + lo: both parts go back to tensor memory as A operands of the last-layer MMAs
        uint32_t lo_bits[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float hi = rna_tf32_mask(sv[i]);
          sv_bits[i] = __float_as_uint(hi);
          lo_bits[i] = __float_as_uint(rna_tf32_mask(sv[i] - hi));
        }
        tmem_st16(t1 + ch * 16, sv_bits);
        tmem_st16(slo + lane_field + ch * 16, lo_bits);
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float4* wrow = reinterpret_cast<const float4*>(Wk + (size_t)(ch * 16 + i) * 32);
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 w4 = wrow[q];
            lacc[q * 4 + 0] = fmaf(sv[i], w4.x, lacc[q * 4 + 0]);
            lacc[q * 4 + 1] = fmaf(sv[i], w4.y, lacc[q * 4 + 1]);
            lacc[q * 4 + 2] = fmaf(sv[i], w4.z, lacc[q * 4 + 2]);
            lacc[q * 4 + 3] = fmaf(sv[i], w4.w, lacc[q * 4 + 3]);
          }
        }
      }
      if (valid) {
        const int nmax = min(16, N - ch * 16);
        if (p.dist != nullptr) {
          float* dst = p.dist + ((size_t)b * N + ch * 16) * HW + pix;
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (i < nmax) dst[(size_t)i * HW] = dv[i];
        }
        if (p.sim != nullptr) {
          float* dst = p.sim + (size_t)m * N + ch * 16;
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (i < nmax) dst[i] = (!X3 && act_log) ? LN2 * sv[i] : sv[i];
        }
        if (lab_ok && ((cmask[ch] >> lab) & 1u)) {
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (pcs[ch * 16 + i] == lab) p.own_dist[(size_t)m * JMAX + prs[ch * 16 + i]] = dv[i];
        }
      }
    };
    {
      // TMEM loads run one chunk ahead of the arithmetic
      uint32_t ra[16], rb[16];
      tmem_ld16(t1, ra);
      for (int ch = 0; ch < nch; ch += 2) {
        tmem_ld_wait();
        if (ch + 1 < nch) tmem_ld16(t1 + (ch + 1) * 16, rb);
        chunk(ch, ra);
        if (ch + 1 < nch) {
          tmem_ld_wait();
          if (ch + 2 < nch) tmem_ld16(t1 + (ch + 2) * 16, ra);
          chunk(ch + 1, rb);
        }
      }
    }
    if (NWG == 2) {
      // hand the token over if the other group still has an epilogue that will wait for it
      const bool other_waits = wg == 0 ? (it + 1 <= n_other) : (it + 2 <= n_other);
      if (other_waits) named_bar_arrive(3 + (1 - wg), 256);
    }

    ADLM_PHASE(ph); ++ph;                      // +5 epilogue 1 done
    float lg[32];
    if (!X3 || tcl) {
      // ---- MMA2: logits = S . W^T with S read from TMEM (3xTF32: Sh.Wh + Sh.Wl + Sl.Wh) ----
      tmem_st_wait();
      fence_before();
      named_bar_sync(1 + wg, 128);
      if (wt == 0) {
        fence_after();
        const uint32_t wb = smem_u32(Wk), wbl = smem_u32(Wk_lo);
        for (int ks = 0; ks < Npad / 8; ++ks) {
          mma_ts(d2, d1 + ks * 8, make_desc(wb + ks * 2 * Cpad * 16, Cpad * 16, 128), idesc2, ks > 0 ? 1u : 0u);
          if (tcl) {
            mma_ts(d2, d1 + ks * 8, make_desc(wbl + ks * 2 * Cpad * 16, Cpad * 16, 128), idesc2, 1u);
            mma_ts(d2, slo + ks * 8, make_desc(wb + ks * 2 * Cpad * 16, Cpad * 16, 128), idesc2, 1u);
          }
        }
        mma_commit(bar_d2);
      }
      mbar_wait(bar_d2, parity);
      fence_after();
      uint32_t r[16];
      tmem_ld16(d2 + lane_field, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 16; ++i) lg[i] = __uint_as_float(r[i]);
      if (Cpad > 16) {
        tmem_ld16(d2 + lane_field + 16, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) lg[16 + i] = __uint_as_float(r[i]);
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) lg[16 + i] = 0.f;
      }
    } else {
#pragma unroll
      for (int c = 0; c < 32; ++c) lg[c] = lacc[X3 ? c : 0];
    }

    ADLM_PHASE(ph); ++ph;                      // +6 MMA2 done, logits in registers
    // ---- epilogue 2: logits + argmax.  The tile's logits are contiguous in HBM ([128 px][C]); they are
    // transposed through the (now idle) A-operand buffer so that the stores are full 16-byte coalesced ----
    {
      float* stg = Ahi;                        // 128 * C floats <= KP * 128 floats
      int best = 0;
      float bv = lg[0];
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        if (c < C) {
          stg[wt * C + c] = lg[c];
          if (c > 0 && lg[c] > bv) { bv = lg[c]; best = c; }
        }
      }
      if (valid && p.argmax != nullptr) p.argmax[m] = best;
      named_bar_sync(1 + wg, 128);
      const int m0 = tile * TILE;
      const int npx = min(TILE, Mi - m0);
      float* dst = p.logits + (size_t)m0 * C;
      const int cnt = npx * C;
      if ((cnt & 3) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {
        for (int i = wt; i < (cnt >> 2); i += 128)
          reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(stg)[i];
      } else {
        for (int i = wt; i < cnt; i += 128) dst[i] = stg[i];
      }
      named_bar_sync(1 + wg, 128);             // staging buffer is the next tile's operand buffer
    }
    // TMEM reads of this tile are complete (wait::ld); order them before the next tile's MMAs
    fence_before();
    ADLM_PHASE(ph); ++ph;                      // +7 stored
  }

  __syncthreads();
  ADLM_PHASE(63);
  if (tid < 32) tmem_dealloc(tmem_base, X3 ? 512 : NWG * 256);
}

int g_num_sms = 0;
int num_sms() {
  if (g_num_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  return g_num_sms;
}

template <int KP, bool X3>
int launch_fwd(const TcFwdParams& prm_in, cudaStream_t st) {
  using L = FwdSmem<KP, X3>;
  TcFwdParams prm = prm_in;
  // 3xTF32: the last layer runs on the tensor cores when its doubled operand fits shared memory and the low parts
  // of the similarities fit tensor memory (2 Npad + Cpad <= 512 columns); ADLM_X3_FMA_LAST_LAYER=1 keeps the fp32 FMAs
  prm.x3_tensor_last = (X3 && L::total(prm.Npad, prm.Cpad, true) + 64 <= 227 * 1024 && 2 * prm.Npad + prm.Cpad <= 512 &&
                        getenv("ADLM_X3_FMA_LAST_LAYER") == nullptr) ? 1 : 0;
  const size_t bytes = L::total(prm.Npad, prm.Cpad, prm.x3_tensor_last != 0) + 64;
  if (bytes > 227 * 1024) {
    set_error("tc_head_fwd: %zu bytes of shared memory needed (> 227 KB) for K=%d N=%d", bytes, prm.K, prm.N);
    return ADLM_E_UNSUPPORTED;
  }
  ADLM_CUDA(cudaFuncSetAttribute(k_tc_fwd<KP, X3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  const int grid = prm.num_tiles < num_sms() ? prm.num_tiles : num_sms();
  KernelTimer timer(ADLM_K_HEAD_FWD, st);
  k_tc_fwd<KP, X3><<<grid, L::NWG * 128, bytes, st>>>(prm);
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

}  // namespace

int tc_head_fwd(const HeadShape& s, int prec_mode, const float* z, const float* P, const float* W,
                float* logits, float* dist, float* sim, int32_t* argmax, const StepExtras& ex,
                cudaStream_t st) {
  ADLM_CHECK_ARG(logits != nullptr || (ex.ce_accum != nullptr && head_fwd_fuses_ce(s, prec_mode)), "logits must be non-null");
  ADLM_CHECK_SUPPORTED(s.K <= 128, "tensor-core path needs K <= 128 (use precision fp32)");
  ADLM_CHECK_SUPPORTED(s.N <= 224, "tensor-core path needs N <= 224 (use precision fp32)");
  ADLM_CHECK_SUPPORTED(s.C <= 32, "tensor-core path needs C <= 32 (use precision fp32)");
  TcFwdParams prm;
  prm.z = z; prm.P = P; prm.W = W;
  prm.logits = logits; prm.dist = dist; prm.sim = sim; prm.argmax = argmax;
  prm.target = ex.target; prm.proto_class = ex.proto_class; prm.proto_rank = ex.proto_rank;
  prm.own_dist = ex.own_dist;
  prm.ce_counts = ex.ce_counts; prm.dlogits = ex.ce_dlogits; prm.ce_accum = ex.ce_accum;
  prm.ce_dp = ex.ce_dp; prm.ce_dp_lay = ex.ce_dp_lay; prm.ce_counts_out = ex.ce_counts_out;
  prm.ce_scale = ex.ce_scale; prm.ignore_void = ex.ignore_void;
  prm.z_out = ex.z_out;
  prm.after_prepare = ex.after_prepare;
  if (ex.z_out != nullptr && (prec_mode == ADLM_PREC_TF32X3 || s.K > 64)) {
    set_error("tc_head_fwd: the fused sigmoid prologue needs ADLM_PREC_TF32 and K <= 64");
    return ADLM_E_UNSUPPORTED;
  }
  prm.B = s.B; prm.K = s.K; prm.HW = s.HW; prm.N = s.N; prm.C = s.C;
  prm.eps = s.eps; prm.act_mode = s.act_mode;
  prm.Npad = (s.N + 15) & ~15;
  prm.Cpad = (s.C + 15) & ~15;
  prm.M = (long long)s.B * s.HW;
  prm.num_tiles = (int)((prm.M + TILE - 1) / TILE);
  const bool x3 = prec_mode == ADLM_PREC_TF32X3;
  if (!x3 && s.K <= 64) return tc_head_fwd_ws(prm, st);     // warp-specialised pipeline (tc_head_ws.cu)
  const int kp = s.K <= 32 ? 32 : (s.K <= 64 ? 64 : 128);
  switch (kp) {
    case 32: return x3 ? launch_fwd<32, true>(prm, st) : launch_fwd<32, false>(prm, st);
    case 64: return x3 ? launch_fwd<64, true>(prm, st) : launch_fwd<64, false>(prm, st);
    default: return x3 ? launch_fwd<128, true>(prm, st) : launch_fwd<128, false>(prm, st);
  }
}

#ifdef ADLM_PHASE_CLOCKS
extern "C" ADLM_API int adlm_debug_phase_clocks(long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, g_phase, sizeof(long long) * 128) == cudaSuccess ? 0 : -3;
}
#endif

}  // namespace adlm
