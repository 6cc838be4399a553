// ADLM_PREC_TF32 backward of the prototype head on the tensor cores: two kernels, both recompute the
// distances on chip (nothing of size M x N is read from or written to HBM).
//
//   G[m][n] = ((dlogits W)[m][n] + dsim[m][n]) * ds/dd(d[m][n]) + ddist[m][n] + kld_grad[m][n], zero where relu clipped
//   dZ = 2 (rowsum(G) zc - G Pc)      k_tc_bwd_dz   lanes = pixels;  G goes back into TMEM and is the A operand
//                                                   of the G.P contraction (.ts MMA)
//   dP = 2 (colsum(G) Pc - G^T Zc)    k_tc_bwd_dp   lanes = prototypes (the transposed problem D^T = Pc Zc^T), so that
//   dW = dlogits^T S                                G and S are again TMEM A operands of the two pixel contractions;
//                                                   accumulators stay in TMEM for all tiles of the CTA, per-CTA
//                                                   partials are summed by k_reduce_partials
// (zc = z - 1/2, Pc = P - 1/2: both formulas are invariant under the common shift.)
// A contraction over the TMEM lane index is impossible (lanes are the M index of an A operand), and MN-major
// tf32 shared-memory operands do not work on this part, hence one kernel per orientation.
#include <stdlib.h>

#include "tc_head.cuh"
#include "tc_common.cuh"

namespace adlm {

namespace {

using namespace tc;

constexpr int TILE = 128;
constexpr int JMAX = ADLM_MAX_PROTOS_PER_CLASS;
constexpr float LN2 = 0.6931471805599453f;
constexpr float CENTRE = 0.5f;

// -DADLM_PHASE_CLOCKS: CTA 0 of k_tc_bwd_dp records clock64() per role (0 builder, 1 MMA issuer, 2 epilogue warp 0)
#ifdef ADLM_PHASE_CLOCKS
__device__ long long g_dp_phase[3][64];
#define DP_PHASE(role, i)                                                                           \
  do {                                                                                              \
    if (blockIdx.x == 0 && (threadIdx.x & 31) == 0 && (i) < 64) g_dp_phase[(role)][(i)] = clock64(); \
  } while (0)
// ... and of k_tc_bwd_dz per warpgroup (thread 0 of each): tile start, operands written, D/Gl done (h = 0), epilogue
// done, D/Gl done (h = 1), epilogue done, dZ done, stored — 8 stamps per tile
__device__ long long g_dz_phase[2][64];
#define DZ_PHASE(i)                                                                          \
  do {                                                                                       \
    if (blockIdx.x == 0 && wt == 0 && (i) < 64) g_dz_phase[wg][(i)] = clock64();             \
  } while (0)
#else
#define DP_PHASE(role, i)
#define DZ_PHASE(i)
#endif

struct TcBwdParams {
  const float* z;
  const float* P;
  const float* W;
  const float* dlogits;   // [M, C] | null
  const float* ddist;     // [B, N, HW] | null
  const float* dsim;      // [M, N] | null
  const int32_t* target;  // [M] | null (KLD term)
  const int32_t* proto_class;
  const int32_t* proto_rank;
  const float* own_grad;  // [M, JMAX] | null
  float* dz;              // [B, K, HW]
  int dz_sigmoid;         // dz <- dz * z (1 - z): gradient wrt the pre-sigmoid features
  float* partials;        // [grid][Npad2*KP + Npad2*32]
  int B, K, HW, N, C;
  float eps;
  int act_mode;
  int Npad, C8;
  long long M;
  int num_tiles;
  int exp_flags;          // timing aid of the fused backward (ADLM_FUSED_EXP; results wrong): 1 no dZ write-out, 2 no dZ MMAs, 4 no G shared-memory stores
};

// ------------------------------------------------------------------------------------------------ dZ
// smem: Pk [KP/4][Npad][4] | Pt [Npad/4][KP][4] | Wt [C8/4][Npad][4] | per WG: Zk [KP/4][128][4], dLk [C8/4][128][4]
template <int KP>
struct DzSmem {
  static __host__ __device__ size_t pk() { return 0; }
  static __host__ __device__ size_t pt(int Npad) { return (size_t)KP * Npad * 4; }
  static __host__ __device__ size_t wt(int Npad) { return 2 * (size_t)KP * Npad * 4; }
  static __host__ __device__ size_t zk(int wg, int Npad, int C8) {
    return wt(Npad) + (size_t)C8 * Npad * 4 + (size_t)wg * ((size_t)KP * TILE * 4 + (size_t)C8 * TILE * 4);
  }
  static __host__ __device__ size_t dlk(int wg, int Npad, int C8) { return zk(wg, Npad, C8) + (size_t)KP * TILE * 4; }
  static __host__ __device__ size_t misc(int Npad, int C8) { return zk(2, Npad, C8); }
  static __host__ __device__ size_t total(int Npad, int C8) { return misc(Npad, C8) + (size_t)Npad * 12 + 256; }
};

template <int KP, bool GENERIC>
__global__ void __launch_bounds__(256, 1) k_tc_bwd_dz(const TcBwdParams p) {
  using L = DzSmem<KP>;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int Npad = p.Npad, C8 = p.C8, K = p.K, N = p.N, C = p.C, HW = p.HW, NH = Npad / 2;
  float* Pk = reinterpret_cast<float*>(smem + L::pk());
  float* Pt = reinterpret_cast<float*>(smem + L::pt(Npad));
  float* Wt = reinterpret_cast<float*>(smem + L::wt(Npad));
  uint8_t* misc = smem + L::misc(Npad, C8);
  float* p2s = reinterpret_cast<float*>(misc);
  int* pcs = reinterpret_cast<int*>(misc + (size_t)Npad * 4);
  int* prs = reinterpret_cast<int*>(misc + (size_t)Npad * 8);
  uint64_t* bars = reinterpret_cast<uint64_t*>(misc + (size_t)Npad * 12);
  uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(bars + 4);
  uint32_t* cmask = reinterpret_cast<uint32_t*>(misc + (size_t)Npad * 12 + 64);   // class bits per 16-prototype chunk

  const int tid = threadIdx.x, wg = tid >> 7, wt = tid & 127, warp_in_wg = wt >> 5;
  const bool kld = p.own_grad != nullptr;

  if (tid < 32) tmem_alloc(smem_u32(tmem_base_s), 512);
  if (tid == 32) {
    for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bars[i]), 1);
    mbar_init_fence();
  }
  // features of the first tile travel while the operands are staged
  const int Mi = (int)p.M;
  const int stride = (int)gridDim.x * 2;
  int tile = (int)blockIdx.x + wg * (int)gridDim.x;
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

  // Pk: rows = prototypes [KP/4][Npad][4]; Wt: rows = prototypes, K index = class [C8/4][Npad][4];
  // Pt: the same rounded prototypes transposed (rows k, K index n) [Npad/4][KP][4]
  {
    // raw copies land in the two warpgroups' pixel-operand buffers (unused until the first tile is converted)
    float* praw = reinterpret_cast<float*>(smem + L::zk(0, Npad, C8));
    float* wraw = praw + (size_t)N * (KP + 4);
    const size_t scratch = 2 * ((size_t)KP * TILE * 4 + (size_t)C8 * TILE * 4);
    if (raw_stage_ok(p.P, p.W, K, KP, C * N) && ((size_t)N * (KP + 4) + (size_t)C * N) * 4 <= scratch) {
      raw_copy_async(praw, wraw, p.P, p.W, N, K, C * N, tid, 256);
      raw_copy_wait();
      __syncthreads();
      pk_from_raw<KP>(Pk, praw, N, Npad, CENTRE, tid, 256);
      wt_from_raw(Wt, wraw, C, N, C8, Npad, 1.f, p.dlogits != nullptr, tid, 256);
      pt_from_raw<KP>(Pt, praw, N, Npad, CENTRE, tid, 256);
    } else {
      stage_rows_kmajor<KP, false>(Pk, Pk, p.P, N, K, Npad, CENTRE, tid, 256);
      stage_wt_kmajor(Wt, p.W, C, N, C8, Npad, 1.f, p.dlogits != nullptr, tid, 256);
      stage_pt_kmajor<KP>(Pt, p.P, N, K, Npad, CENTRE, tid, 256);
    }
  }
  __syncthreads();
  for (int n = tid; n < Npad; n += 256) {
    float a = 0.f;
    for (int k4 = 0; k4 < KP / 4; ++k4) {
      const float4 v = *reinterpret_cast<const float4*>(Pk + ((size_t)k4 * Npad + n) * 4);
      a = fmaf(v.x, v.x, a); a = fmaf(v.y, v.y, a); a = fmaf(v.z, v.z, a); a = fmaf(v.w, v.w, a);
    }
    p2s[n] = a;
    const int rk = (kld && n < N) ? p.proto_rank[n] : -1;    // < 0: beyond ADLM_MAX_PROTOS_PER_CLASS
    pcs[n] = rk >= 0 ? p.proto_class[n] : -1;
    prs[n] = rk >= 0 ? rk : 0;
  }
  __syncthreads();
  for (int c = tid; c < Npad / 16; c += 256) {
    uint32_t mk = 0u;
    for (int i = 0; i < 16; ++i) {
      const int pc = pcs[c * 16 + i];
      if (pc >= 0 && pc < 32) mk |= 1u << pc;
    }
    cmask[c] = mk;
  }
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  // Inside the step this grid is launched while k_kld still runs.  k_kld's output (the own-class gradients) is first
  // touched after the first tile's operands are written (`waited` below), so staging and that conversion run under its
  // tail; the dP/dW kernel, which reads nothing this kernel writes, may in turn start as soon as CTAs of this grid
  // leave their SMs — CTAs with one tile less (529 tiles over 148 CTAs) hand over early.  The dependents are released
  // only AFTER this grid's own wait: the dP/dW kernel reads k_kld's output too and has no wait of its own (a
  // programmatic launch orders a grid after its immediate predecessor only), so it must not start before k_kld is
  // known to be complete (test_step_many_tiles_per_cta_equals_sum_of_single_image_steps caught the other order).
  bool waited = false;

  const uint32_t tmem_base = *tmem_base_s;
  const uint32_t tD = tmem_base + wg * 256;      // D1 half   [128 x NH]
  const uint32_t tG = tD + NH;                   // Gl half   [128 x NH]
  const uint32_t tZ = tD + 2 * NH;               // dZ        [128 x KP]
  const uint32_t lane_field = (uint32_t)(warp_in_wg * 32) << 16;
  float* Zk = reinterpret_cast<float*>(smem + L::zk(wg, Npad, C8));
  float* dLk = reinterpret_cast<float*>(smem + L::dlk(wg, Npad, C8));
  const uint32_t bar_a = smem_u32(&bars[wg * 2]), bar_z = smem_u32(&bars[wg * 2 + 1]);
  const uint32_t idesc_h = make_idesc_tf32(TILE, NH), idesc_z = make_idesc_tf32(TILE, KP);
  const int ksteps = (K + 7) >> 3;

  uint32_t ph_a = 0, ph_z = 0;
  const float eps = p.eps, epsm1 = p.eps - 1.f;
  const bool act_log = p.act_mode == ADLM_ACT_LOG;

  int dzph = 1;
  DZ_PHASE(0);
  for (; tile < p.num_tiles; tile += stride) {
    const int m = tile * TILE + wt;
    const bool valid = m < Mi;
    const int b = valid ? m / HW : 0;
    const int pix = valid ? m - b * HW : 0;
    DZ_PHASE(dzph); ++dzph;                  // tile start

    // ---- operands of this tile: Zk (centred, rounded), dLk ----
    float x2 = 0.f;
#pragma unroll
    for (int k4 = 0; k4 < KP / 4; ++k4) {
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[j] = rna_tf32_mask(zr[k4 * 4 + j] - CENTRE);
        x2 = fmaf(v[j], v[j], x2);
      }
      *reinterpret_cast<float4*>(Zk + ((size_t)k4 * TILE + wt) * 4) = make_float4(v[0], v[1], v[2], v[3]);
    }
    for (int c4 = 0; c4 < C8 / 4; ++c4) {
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int c = c4 * 4 + j;
        v[j] = (valid && c < C && p.dlogits != nullptr) ? rna_tf32_fast(__ldg(p.dlogits + (size_t)m * C + c)) : 0.f;
      }
      *reinterpret_cast<float4*>(dLk + ((size_t)c4 * TILE + wt) * 4) = make_float4(v[0], v[1], v[2], v[3]);
    }
    int lab = -1;
    if (kld && valid) lab = p.target[m];
    const bool lab_ok = lab >= 0 && lab < 32;
    fence_async_smem();
    fence_before();
    named_bar_sync(1 + wg, 128);
    prefetch(tile + stride);
    if (!waited) {                           // no-ops unless launched with programmatic stream serialisation
      asm volatile("griddepcontrol.wait;" ::: "memory");
      asm volatile("griddepcontrol.launch_dependents;");
      waited = true;
    }
    // pull the pixel's 64-byte line of own-class KLD gradients into L1 before the epilogue needs it (after the wait:
    // a line fetched while k_kld still writes it would sit stale in L1)
    if (kld && valid) asm volatile("prefetch.global.L1 [%0];" ::"l"(p.own_grad + (size_t)m * JMAX));
    DZ_PHASE(dzph); ++dzph;                  // operands written, next tile's loads issued

    float rs = 0.f;
    // (rolled: the tile loop's body is 64 KB of SASS when both halves are expanded, more than the instruction cache
    // holds — ncu attributes ~19 % of the warps' time to instruction fetch, profiles/r2_ncu_full_step_tf32.txt)
#pragma unroll 1
    for (int h = 0; h < 2; ++h) {
      // MMAs are issued by the warpgroup's first warp as a whole (one elected lane per instruction): from a single
      // thread every tcgen05.mma costs ~120 cycles of descriptor shuffling into uniform registers (tc_common.cuh)
      if (warp_in_wg == 0) {
        fence_after();
        const uint32_t a0 = smem_u32(Zk), b0 = smem_u32(Pk) + h * NH * 16;
        for (int ks = 0; ks < ksteps; ++ks)
          mma_ss_w(tD, make_desc(a0 + ks * 2 * TILE * 16, TILE * 16, 128), make_desc(b0 + ks * 2 * Npad * 16, Npad * 16, 128),
                   idesc_h, ks > 0 ? 1u : 0u);
        const uint32_t l0 = smem_u32(dLk), w0 = smem_u32(Wt) + h * NH * 16;
        for (int ks = 0; ks < C8 / 8; ++ks)
          mma_ss_w(tG, make_desc(l0 + ks * 2 * TILE * 16, TILE * 16, 128), make_desc(w0 + ks * 2 * Npad * 16, Npad * 16, 128),
                   idesc_h, ks > 0 ? 1u : 0u);
        mma_commit_w(bar_a);
      }
      mbar_wait(bar_a, ph_a);
      ph_a ^= 1u;
      fence_after();
      DZ_PHASE(dzph); ++dzph;                // D / Gl of this half done
      for (int ch = 0; ch < NH / 16; ++ch) {
        uint32_t ra[16], rg[16], out[16];
        tmem_ld16(tD + lane_field + ch * 16, ra);
        tmem_ld16(tG + lane_field + ch * 16, rg);
        const int n0 = h * NH + ch * 16;
        float xp[16];
#pragma unroll
        for (int v4 = 0; v4 < 4; ++v4) {
          const float4 q = *reinterpret_cast<const float4*>(p2s + n0 + v4 * 4);
          xp[v4 * 4] = x2 + q.x; xp[v4 * 4 + 1] = x2 + q.y; xp[v4 * 4 + 2] = x2 + q.z; xp[v4 * 4 + 3] = x2 + q.w;
        }
        tmem_ld_wait();
        float g[16];
        uint32_t pos = 0u;
        // padding columns and out-of-range pixels need no masking: their dlogits.W is exactly 0
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float pre = fmaf(-2.f, __uint_as_float(ra[i]), xp[i]);
          const float d = fmaxf(pre, 0.f);
          float gv = __uint_as_float(rg[i]);
          if (GENERIC) {
            const int n = n0 + i;
            if (p.dsim != nullptr && valid && n < N) gv += __ldg(p.dsim + (size_t)m * N + n);
          }
          gv *= act_log ? epsm1 * rcp_approx((d + 1.f) * (d + eps)) : -1.f;
          if (GENERIC) {
            const int n = n0 + i;
            if (p.ddist != nullptr && valid && n < N) gv += __ldg(p.ddist + ((size_t)b * N + n) * HW + pix);
          }
          pos |= (pre > 0.f ? 1u : 0u) << i;
          g[i] = gv;
        }
        if (lab_ok && ((cmask[n0 >> 4] >> lab) & 1u)) {     // KLD term: only the pixel's own-class prototypes
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (pcs[n0 + i] == lab) g[i] += __ldg(p.own_grad + (size_t)m * JMAX + prs[n0 + i]);
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float gv = ((pos >> i) & 1u) ? g[i] : 0.f;
          rs += gv;
          out[i] = __float_as_uint(rna_tf32_fast(gv));
        }
        tmem_st16(tD + lane_field + ch * 16, out);
      }
      tmem_st_wait();
      fence_before();
      named_bar_sync(1 + wg, 128);
      DZ_PHASE(dzph); ++dzph;                // gradient epilogue of this half done
      if (warp_in_wg == 0) {
        fence_after();
        const uint32_t pt0 = smem_u32(Pt) + (h * NH / 4) * KP * 16;
        for (int ks = 0; ks < NH / 8; ++ks)
          mma_ts_w(tZ, tD + ks * 8, make_desc(pt0 + ks * 2 * KP * 16, KP * 16, 128), idesc_z, (h > 0 || ks > 0) ? 1u : 0u);
        if (h == 1) mma_commit_w(bar_z);
      }
    }
    mbar_wait(bar_z, ph_z);
    ph_z ^= 1u;
    fence_after();
    DZ_PHASE(dzph); ++dzph;                  // dZ accumulator complete
    // ---- dz = 2 (rowsum(G) zc - G Pc); zc re-read from the operand tile (tf32-rounded, 2^-12 relative) ----
    float* dst = p.dz + ((size_t)b * K) * HW + pix;
#pragma unroll 1
    for (int ch = 0; ch < KP / 16; ++ch) {
      uint32_t r[16];
      tmem_ld16(tZ + lane_field + ch * 16, r);
      float zc[16];
#pragma unroll
      for (int v4 = 0; v4 < 4; ++v4) {
        const float4 q = *reinterpret_cast<const float4*>(Zk + ((size_t)(ch * 4 + v4) * TILE + wt) * 4);
        zc[v4 * 4] = q.x; zc[v4 * 4 + 1] = q.y; zc[v4 * 4 + 2] = q.z; zc[v4 * 4 + 3] = q.w;
      }
      tmem_ld_wait();
      if (valid) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int k = ch * 16 + i;
          if (k < K) {
            float gz = 2.f * fmaf(rs, zc[i], -__uint_as_float(r[i]));
            if (p.dz_sigmoid) {                       // z = zc + 1/2 (tf32-rounded): z (1 - z) = 1/4 - zc^2
              gz *= fmaf(-zc[i], zc[i], 0.25f);
            }
            dst[(size_t)k * HW] = gz;
          }
        }
      }
    }
    fence_before();
    DZ_PHASE(dzph); ++dzph;                  // stored
  }
  DZ_PHASE(63);
  __syncthreads();
  if (tid < 32) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------- dP, dW
// lanes = prototypes.  Roles: warps 0-7 epilogue (TMEM lane n = 32*(warp%4)+lane, pixel columns split in two
// halves by warp/4), warps 8-11 build the pixel-side operands (thread = pixel), warp 12 issues the MMAs.
// smem: ZtB (128B-swizzled [KP][128 px]) | dLtB (swizzled [32][128 px]) | PkA [KP/4][Npad2][4] |
//       WtA [C8/4][Npad2][4] | ZkB [KP/4][128][4] | dLkB [C8/4][128][4] | x2, label, image, pixel per column
constexpr int DP_THREADS = 416;

template <int KP>
struct DpSmem {
  static __host__ __device__ size_t zt() { return 0; }
  static __host__ __device__ size_t dlt() { return (size_t)KP * TILE * 4; }
  static __host__ __device__ size_t pk() { return dlt() + 32 * TILE * 4; }
  static __host__ __device__ size_t wt(int Npad2) { return pk() + (size_t)KP * Npad2 * 4; }
  static __host__ __device__ size_t zk(int Npad2, int C8) { return wt(Npad2) + (size_t)C8 * Npad2 * 4; }
  static __host__ __device__ size_t dlk(int Npad2, int C8) { return zk(Npad2, C8) + (size_t)KP * TILE * 4; }
  static __host__ __device__ size_t misc(int Npad2, int C8) { return dlk(Npad2, C8) + (size_t)C8 * TILE * 4; }
  // per-tile metadata in two slots (tile parity): x2 [2][128] f32, img [2][128] i32, pix [2][128] i32,
  // class pixel masks [2][32][4] u32; then cs [2][2][128] f32, barriers + TMEM base (128 B)
  // ... + own-class KLD gradients of the tile [2][128 px][16] f32 (cp.async by the builder warps)
  static constexpr size_t MISC_BYTES = 3 * 1024 + 1024 + 2048 + 128;
  static __host__ __device__ size_t ogs(int Npad2, int C8) { return misc(Npad2, C8) + MISC_BYTES; }
  static __host__ __device__ size_t total(int Npad2, int C8) { return ogs(Npad2, C8) + 2 * 128 * 16 * 4; }
};

template <int KP, int NB, bool GENERIC>
// 13 warps: the sub-partition that hosts four of them caps the kernel at 16384 / (4 * 32) = 128 registers
__global__ void __launch_bounds__(DP_THREADS, 1) k_tc_bwd_dp(const TcBwdParams p) {
  using L = DpSmem<KP>;
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr int Npad2 = NB * 128;
  const int C8 = p.C8, K = p.K, N = p.N, C = p.C, HW = p.HW;
  float* ZtB = reinterpret_cast<float*>(smem + L::zt());
  float* dLtB = reinterpret_cast<float*>(smem + L::dlt());
  float* PkA = reinterpret_cast<float*>(smem + L::pk());
  float* WtA = reinterpret_cast<float*>(smem + L::wt(Npad2));
  float* ZkB = reinterpret_cast<float*>(smem + L::zk(Npad2, C8));
  float* dLkB = reinterpret_cast<float*>(smem + L::dlk(Npad2, C8));
  uint8_t* misc = smem + L::misc(Npad2, C8);
  float* x2s = reinterpret_cast<float*>(misc);                      // [slot][128]
  int* imgs = reinterpret_cast<int*>(misc + 1024);                   // [slot][128]
  int* pixs = reinterpret_cast<int*>(misc + 2048);                   // [slot][128]
  uint32_t* clsm = reinterpret_cast<uint32_t*>(misc + 3072);         // [slot][32 classes][4 words]: pixels of the tile
  float* css = reinterpret_cast<float*>(misc + 4096);                // [half][blk][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(misc + 4096 + 2048);
  uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(bars + 8);     // eight barriers, then the TMEM base
  float* ogs = reinterpret_cast<float*>(smem + L::ogs(Npad2, C8));   // [slot][128 px][16]: dL/d(own distances)

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool kld = p.own_grad != nullptr;
  const bool has_dl = p.dlogits != nullptr;
  if (warp == 0) DP_PHASE(2, 60);

  if (warp == 12) tmem_alloc(smem_u32(tmem_base_s), 512);
  if (tid == 0) {
    mbar_init(smem_u32(&bars[0]), 128);   // full_k: K-major pixel operands (distance MMAs) + tile metadata written
    mbar_init(smem_u32(&bars[1]), 1);     // free: every MMA of the tile done
    mbar_init(smem_u32(&bars[2]), 1);     // acc[b]: D^T and Gl^T of a unit ready in TMEM buffer b
    mbar_init(smem_u32(&bars[3]), 1);
    mbar_init(smem_u32(&bars[4]), 256);   // gs[b]: G and S of the unit written back to TMEM buffer b
    mbar_init(smem_u32(&bars[5]), 256);
    mbar_init(smem_u32(&bars[6]), 128);   // full_t: transposed pixel operands (accumulate MMAs) written
    mbar_init(smem_u32(&bars[7]), 1);     // k_free: the tile's last distance MMA has read ZkB / dLkB (one completion per tile)
    mbar_init_fence();
  }
  // the builder warps fetch their first tile before the one-time operand staging
  const int iters = (p.num_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1;
  const bool is_builder = warp >= 8 && warp < 12;
  const int cw = tid - 256;
  float zr[KP];
  float dlr[32];
  const int Mi = (int)p.M;
  auto prefetch = [&](int it) {
    const int m = ((int)blockIdx.x + it * (int)gridDim.x) * TILE + cw;
    if (it < iters && m < Mi) {
      const int b = m / HW;
      const float* src = p.z + ((size_t)b * K) * HW + (m - b * HW);
      if (K == KP) {
#pragma unroll
        for (int k = 0; k < KP; ++k) { zr[k] = __ldg(src); src += HW; }
      } else {
#pragma unroll
        for (int k = 0; k < KP; ++k) zr[k] = (k < K) ? __ldg(src + (size_t)k * HW) : CENTRE;
      }
      const float* dl = p.dlogits + (size_t)m * C;
#pragma unroll
      for (int c = 0; c < 32; ++c) dlr[c] = (has_dl && c < C) ? __ldg(dl + c) : 0.f;
    } else {
#pragma unroll
      for (int k = 0; k < KP; ++k) zr[k] = CENTRE;
#pragma unroll
      for (int c = 0; c < 32; ++c) dlr[c] = 0.f;
    }
  };
  if (warp < 8) {
    // one-time operands, staged by the eight epilogue warps only: the builder warps have their loads in flight
    // and the MMA warp is allocating TMEM
    constexpr int PB = Npad2 * (KP / 4) / 256, WB = 4 * NB;      // C8 <= 32: (C8 / 4) * Npad2 <= WB * 256
    if (K == KP && (reinterpret_cast<uintptr_t>(p.P) & 15u) == 0 && PB <= 16) {
      stage_rows_and_wt_once<KP, (PB <= 16 ? PB : 1), WB>(PkA, WtA, p.P, p.W, N, C, C8, Npad2, CENTRE, has_dl, tid, 256);
    } else {
      stage_rows_kmajor<KP, false, 4>(PkA, PkA, p.P, N, K, Npad2, CENTRE, tid, 256);
      stage_wt_kmajor(WtA, p.W, C, N, C8, Npad2, 1.f, has_dl, tid, 256);
    }
    smem_zero4(dLtB, 32 * TILE / 4, tid, 256);
  }
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();

  const uint32_t tmem_base = *tmem_base_s;
  if (warp == 8) DP_PHASE(0, 0);
  if (warp == 12) DP_PHASE(1, 0);
  if (warp == 0) DP_PHASE(2, 0);
  // A unit = (prototype block, 64-pixel half of the tile).  Two TMEM buffers of 128 columns (D^T | Gl^T, 64 each)
  // let the tensor pipe produce unit u+1 and consume unit u-1 while the epilogue warps work on unit u.
  constexpr int UW = 64, NU = 2 * NB;
  const uint32_t tP0 = tmem_base + 256, tW0 = tmem_base + 256 + NB * KP;
  auto tDb = [&](int b) { return tmem_base + (uint32_t)b * 128u; };
  // The units of consecutive tiles form one stream.  The builder writes a tile in two steps: the K-major operands
  // and the metadata as soon as the previous tile's last distance MMA has retired (they are consumed two units
  // ahead of the epilogue), the transposed operands once the previous tile's accumulate MMAs are done.
  const uint32_t bar_full = smem_u32(&bars[0]), bar_free = smem_u32(&bars[1]), bar_acc0 = smem_u32(&bars[2]),
                 bar_gs0 = smem_u32(&bars[4]), bar_full_t = smem_u32(&bars[6]), bar_kfree = smem_u32(&bars[7]);

  if (is_builder) {
    prefetch(0);
    // ================= pixel-side operand builder: thread = pixel =================
    for (int it = 0; it < iters; ++it) {
      const int m = ((int)blockIdx.x + it * (int)gridDim.x) * TILE + cw;
      const bool valid = m < Mi;
      const int slot = it & 1;
      uint32_t* clsm_s = clsm + slot * 128;
      float* ogs_s = ogs + slot * 128 * JMAX;
      if (warp == 8) DP_PHASE(0, 1 + it * 8);
      // ---- step K: wait until the last distance MMA of the previous tile has read ZkB / dLkB.  k_free completes
      //      ONCE per tile and its next completion needs this step's full_k, so a parity wait cannot alias however
      //      late this thread arrives.  (The first version waited for the two per-tile completions of acc[1] in
      //      order; a builder held up in the prefetch of the next tile for longer than two epilogue units — HBM
      //      misses at B = 256 on some boxes — found both already done, the wait for the first one then matched the
      //      parity of the NEXT phase and the CTA dead-locked until mbar_wait's watchdog trapped.)
      if (it > 0) mbar_wait(bar_kfree, (uint32_t)(it - 1) & 1u);
      if (warp == 8) DP_PHASE(0, 2 + it * 8);
      clsm_s[cw] = 0u;
      if (kld && valid) {
        // the tile's own-class KLD gradients travel straight into shared memory while the operands are converted
        const float* src = p.own_grad + (size_t)m * JMAX;
        const uint32_t dst = smem_u32(ogs_s + cw * JMAX);
#pragma unroll
        for (int v = 0; v < JMAX / 4; ++v)
          asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst + v * 16), "l"(src + v * 4) : "memory");
      }
      named_bar_sync(2, 128);
      const int mylab = (valid && kld) ? p.target[m] : -1;
      if (mylab >= 0 && mylab < 32) atomicOr(&clsm_s[mylab * 4 + (cw >> 5)], 1u << (cw & 31));
      float x2 = 0.f;
#pragma unroll
      for (int k4 = 0; k4 < KP / 4; ++k4) {
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          v[j] = rna_tf32_mask(zr[k4 * 4 + j] - CENTRE);
          zr[k4 * 4 + j] = v[j];
          x2 = fmaf(v[j], v[j], x2);
        }
        *reinterpret_cast<float4*>(ZkB + ((size_t)k4 * TILE + cw) * 4) = make_float4(v[0], v[1], v[2], v[3]);
      }
#pragma unroll
      for (int c4 = 0; c4 < 8; ++c4) {
        if (c4 * 4 < C8) {
          float v[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            v[j] = rna_tf32_fast(dlr[c4 * 4 + j]);
            dlr[c4 * 4 + j] = v[j];
          }
          *reinterpret_cast<float4*>(dLkB + ((size_t)c4 * TILE + cw) * 4) = make_float4(v[0], v[1], v[2], v[3]);
        }
      }
      x2s[slot * 128 + cw] = x2;
      const int b = valid ? m / HW : 0;
      imgs[slot * 128 + cw] = b;
      pixs[slot * 128 + cw] = valid ? m - b * HW : 0;
      asm volatile("cp.async.wait_all;" ::: "memory");
      fence_async_smem();
      mbar_arrive(bar_full);
      if (warp == 8) DP_PHASE(0, 3 + it * 8);
      // ---- step T: the transposed operands, once the previous tile's accumulate MMAs have retired
      if (it > 0) mbar_wait(bar_free, (uint32_t)(it - 1) & 1u);
#pragma unroll
      for (int k = 0; k < KP; ++k) ZtB[off_sw128(k, cw, KP)] = zr[k];
#pragma unroll
      for (int c = 0; c < 32; ++c)
        if (c < C) dLtB[off_sw128(c, cw, 32)] = dlr[c];
      fence_async_smem();
      mbar_arrive(bar_full_t);
      prefetch(it + 1);
      if (warp == 8) DP_PHASE(0, 4 + it * 8);
    }
  } else if (warp == 12) {
    // ================= MMA issuer (the whole warp, one elected lane per instruction: tc_common.cuh) =================
    {
      const uint32_t idesc_t = make_idesc_tf32(128, UW), idesc_p = make_idesc_tf32(128, KP), idesc_w = make_idesc_tf32(128, 32);
      const int ksteps = (K + 7) >> 3;
      // unit j of a tile: block j >> 1, pixel half j & 1
      auto issue_d = [&](int j) {
        const int blk = j >> 1, h = j & 1, b = j & 1;
        const uint32_t a0 = smem_u32(PkA) + blk * 128 * 16, b0 = smem_u32(ZkB) + h * UW * 16;
        for (int ks = 0; ks < ksteps; ++ks)
          mma_ss_w(tDb(b), make_desc(a0 + ks * 2 * Npad2 * 16, Npad2 * 16, 128), make_desc(b0 + ks * 2 * TILE * 16, TILE * 16, 128),
                 idesc_t, ks > 0 ? 1u : 0u);
        const uint32_t w0 = smem_u32(WtA) + blk * 128 * 16, l0 = smem_u32(dLkB) + h * UW * 16;
        for (int ks = 0; ks < C8 / 8; ++ks)
          mma_ss_w(tDb(b) + UW, make_desc(w0 + ks * 2 * Npad2 * 16, Npad2 * 16, 128), make_desc(l0 + ks * 2 * TILE * 16, TILE * 16, 128),
                 idesc_t, ks > 0 ? 1u : 0u);
        mma_commit_w(bar_acc0 + b * 8);
        if (j == NU - 1) mma_commit_w(bar_kfree);
      };
      uint32_t gcnt[2] = {0u, 0u};
      mbar_wait(bar_full, 0u);
      fence_after();
      issue_d(0);
      issue_d(1);
      for (int it = 0; it < iters; ++it) {
        DP_PHASE(1, 1 + it * 8);
#pragma unroll
        for (int j = 0; j < NU; ++j) {
          const int blk = j >> 1, h = j & 1, b = j & 1;
          mbar_wait(bar_gs0 + b * 8, gcnt[b] & 1u);
          ++gcnt[b];
          if (j == 0) mbar_wait(bar_full_t, (uint32_t)it & 1u);
          fence_after();
          if (j == 0) DP_PHASE(1, 3 + it * 8);
          const uint32_t zt0 = smem_u32(ZtB), lt0 = smem_u32(dLtB);
          for (int k8 = 0; k8 < UW / 8; ++k8) {
            const int ks = h * (UW / 8) + k8;
            const uint32_t acc = (it > 0 || ks > 0) ? 1u : 0u;
            mma_ts_w(tP0 + blk * KP, tDb(b) + k8 * 8, make_desc_sw128(zt0 + (ks >> 2) * KP * 128 + (ks & 3) * 32), idesc_p, acc);
            mma_ts_w(tW0 + blk * 32, tDb(b) + UW + k8 * 8, make_desc_sw128(lt0 + (ks >> 2) * 32 * 128 + (ks & 3) * 32), idesc_w, acc);
          }
          if (j == NU - 1) mma_commit_w(bar_free);
          // the distance MMAs two units ahead go into the buffer the accumulate MMAs above read: issue order
          if (j + 2 < NU) {
            issue_d(j + 2);
          } else if (it + 1 < iters) {
            if (j + 2 - NU == 0) {
              mbar_wait(bar_full, (uint32_t)(it + 1) & 1u);
              fence_after();
            }
            issue_d(j + 2 - NU);
          }
        }
        DP_PHASE(1, 8 + it * 8);
      }
    }
  } else {
    // ================= epilogue: thread = (prototype lane, pixel-column half) =================
    const int half = warp >> 2, lane_n = ((warp & 3) << 5) + lane;
    const uint32_t lane_field = (uint32_t)((warp & 3) * 32) << 16;
    float p2[NB], cs[NB];
    int pcl[NB], prk[NB];
#pragma unroll
    for (int blk = 0; blk < NB; ++blk) {
      const int n = blk * 128 + lane_n;
      float a = 0.f;
      for (int k4 = 0; k4 < KP / 4; ++k4) {
        const float4 v = *reinterpret_cast<const float4*>(PkA + ((size_t)k4 * Npad2 + n) * 4);
        a = fmaf(v.x, v.x, a); a = fmaf(v.y, v.y, a); a = fmaf(v.z, v.z, a); a = fmaf(v.w, v.w, a);
      }
      p2[blk] = a;
      cs[blk] = 0.f;
      pcl[blk] = -1;
      prk[blk] = 0;
      if (kld && n < N) { pcl[blk] = p.proto_class[n]; prk[blk] = p.proto_rank[n]; }
      if (pcl[blk] < 0 || pcl[blk] >= 32 || prk[blk] < 0) { pcl[blk] = -1; prk[blk] = 0; }
    }
    const float eps = p.eps, epsm1 = p.eps - 1.f;
    const bool act_log = p.act_mode == ADLM_ACT_LOG;
    uint32_t acnt[2] = {0u, 0u};
    for (int it = 0; it < iters; ++it) {
      const long long m0 = ((long long)blockIdx.x + (long long)it * gridDim.x) * TILE;
      const int npx = (int)((p.M - m0) < TILE ? (p.M - m0) : TILE);   // valid pixel columns of this tile
      if (warp == 0) DP_PHASE(2, 1 + it * 8);
      mbar_wait(bar_full, (uint32_t)it & 1u);
      const int slot = it & 1;
      const float* x2s_s = x2s + slot * 128;
      const uint32_t* clsm_s = clsm + slot * 128;
      const float* ogs_s = ogs + slot * 128 * JMAX;
      const int* imgs_s = imgs + slot * 128;
      const int* pixs_s = pixs + slot * 128;
#pragma unroll
      for (int j = 0; j < NU; ++j) {
        const int blk = j >> 1, h = j & 1, b = j & 1;
        const int n = blk * 128 + lane_n;
        const uint32_t tD1 = tDb(b), tGl = tDb(b) + UW;
        if (warp == 0 && h == 0) DP_PHASE(2, 2 + it * 8 + blk * 3);
        mbar_wait(bar_acc0 + b * 8, acnt[b] & 1u);
        if (warp == 0 && h == 0) DP_PHASE(2, 3 + it * 8 + blk * 3);
        ++acnt[b];
        fence_after();
        if (blk * 128 + (warp & 3) * 32 < N) {       // warps whose 32 prototypes are all padding skip the math
          for (int ch = 0; ch < 2; ++ch) {
            const int cl = half * 32 + ch * 16;       // column inside the unit's TMEM buffer
            const int c0 = h * UW + cl;               // pixel column of the tile
            uint32_t ra[16], rg[16], og[16], os[16];
            tmem_ld16(tD1 + lane_field + cl, ra);
            tmem_ld16(tGl + lane_field + cl, rg);
            float xp[16];
#pragma unroll
            for (int v4 = 0; v4 < 4; ++v4) {
              const float4 q = *reinterpret_cast<const float4*>(x2s_s + c0 + v4 * 4);
              xp[v4 * 4] = q.x + p2[blk]; xp[v4 * 4 + 1] = q.y + p2[blk]; xp[v4 * 4 + 2] = q.z + p2[blk]; xp[v4 * 4 + 3] = q.w + p2[blk];
            }
            // own-class pixels of this prototype among the 16 columns (KLD term), and the valid columns
            const uint32_t hit = pcl[blk] >= 0 ? ((clsm_s[pcl[blk] * 4 + (c0 >> 5)] >> (c0 & 31)) & 0xffffu) : 0u;
            const int nv = npx - c0;
            const uint32_t vmask = (n < N && nv > 0) ? (nv >= 16 ? 0xffffu : ((1u << nv) - 1u)) : 0u;
            tmem_ld_wait();
            float gvv[16], sv[16];
            uint32_t pos = 0u;
            // out-of-range pixel columns / padding prototypes: dlogits.W is exactly 0 there, only S needs masking
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float pre = fmaf(-2.f, __uint_as_float(ra[i]), xp[i]);
              const float d = fmaxf(pre, 0.f);
              float gv = __uint_as_float(rg[i]);
              if (GENERIC) {
                if (p.dsim != nullptr && ((vmask >> i) & 1u)) gv += __ldg(p.dsim + (size_t)(m0 + c0 + i) * N + n);
              }
              if (act_log) {
                const float u = d + 1.f, v = d + eps;
                sv[i] = lg2_approx(u) - lg2_approx(v);          // log2 units; ln 2 is folded into the flush
                gv *= epsm1 * rcp_approx(u * v);
              } else {
                sv[i] = -d;
                gv = -gv;
              }
              if (GENERIC) {
                if (p.ddist != nullptr && ((vmask >> i) & 1u))
                  gv += __ldg(p.ddist + ((size_t)imgs_s[c0 + i] * N + n) * HW + pixs_s[c0 + i]);
              }
              pos |= (pre > 0.f ? 1u : 0u) << i;
              gvv[i] = gv;
            }
            if (hit) {
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if ((hit >> i) & 1u) gvv[i] += ogs_s[(c0 + i) * JMAX + prk[blk]];
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float gv = ((pos >> i) & 1u) ? gvv[i] : 0.f;
              cs[blk] += gv;
              og[i] = __float_as_uint(rna_tf32_fast(gv));
              os[i] = __float_as_uint(rna_tf32_fast(((vmask >> i) & 1u) ? sv[i] : 0.f));
            }
            tmem_st16(tD1 + lane_field + cl, og);
            tmem_st16(tGl + lane_field + cl, os);
          }
          tmem_st_wait();
        }
        fence_before();
        mbar_arrive(bar_gs0 + b * 8);
        if (warp == 0 && h == 1) DP_PHASE(2, 4 + it * 8 + blk * 3);
      }
    }
    // ---- flush this CTA's partial sums ----
    mbar_wait(bar_free, (uint32_t)(iters - 1) & 1u);
    fence_after();
#pragma unroll
    for (int blk = 0; blk < NB; ++blk) css[(half * 2 + blk) * 128 + lane_n] = cs[blk];
    if (NB == 1) css[(half * 2 + 1) * 128 + lane_n] = 0.f;
    named_bar_sync(1, 256);
    const float s_scale = act_log ? LN2 : 1.f;
    float* part = p.partials + (size_t)blockIdx.x * ((size_t)Npad2 * KP + (size_t)Npad2 * 32);
    // Every MMA has retired (bar_free), so the pixel-operand buffers serve as per-warp transpose scratch:
    // a thread owns one prototype row in TMEM, but the stores should cover whole rows of the partials.
    constexpr int HK = KP / 2, PITCH = HK + 4;                      // 16-byte aligned, conflict-free float4 rows
    static_assert(8 * 32 * PITCH * 4 <= (KP + 32) * TILE * 4, "transpose scratch exceeds ZtB + dLtB");
    float* scr = ZtB + (size_t)warp * 32 * PITCH;
    for (int blk = 0; blk < NB; ++blk) {
      const int n = blk * 128 + lane_n, n_base = blk * 128 + (warp & 3) * 32;
      const float cst = css[blk * 128 + lane_n] + css[(2 + blk) * 128 + lane_n];
      if (n_base < N) {
#pragma unroll
        for (int ch = 0; ch < KP / 32; ++ch) {
          const int k0 = half * HK + ch * 16;
          uint32_t r[16];
          tmem_ld16(tP0 + blk * KP + lane_field + k0, r);
          tmem_ld_wait();
#pragma unroll
          for (int i4 = 0; i4 < 4; ++i4) {
            float o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int k = k0 + i4 * 4 + j;
              const float pc = PkA[(size_t)(k >> 2) * Npad2 * 4 + n * 4 + (k & 3)];
              o[j] = 2.f * (cst * pc - __uint_as_float(r[i4 * 4 + j]));
            }
            *reinterpret_cast<float4*>(scr + lane * PITCH + ch * 16 + i4 * 4) = make_float4(o[0], o[1], o[2], o[3]);
          }
        }
        __syncwarp();
        {
          constexpr int V = HK / 4, RPI = 32 / V;                   // float4s per row, rows per store instruction
          const int col4 = lane % V, rsub = lane / V;
#pragma unroll
          for (int i = 0; i < 32 / RPI; ++i) {
            const int row = i * RPI + rsub;
            if (n_base + row < N)
              *reinterpret_cast<float4*>(part + (size_t)(n_base + row) * KP + half * HK + col4 * 4) =
                  *reinterpret_cast<const float4*>(scr + row * PITCH + col4 * 4);
          }
        }
        __syncwarp();
        uint32_t r[16];
        tmem_ld16(tW0 + blk * 32 + lane_field + half * 16, r);
        tmem_ld_wait();
#pragma unroll
        for (int i4 = 0; i4 < 4; ++i4)
          *reinterpret_cast<float4*>(scr + lane * PITCH + i4 * 4) =
              make_float4(s_scale * __uint_as_float(r[i4 * 4]), s_scale * __uint_as_float(r[i4 * 4 + 1]),
                          s_scale * __uint_as_float(r[i4 * 4 + 2]), s_scale * __uint_as_float(r[i4 * 4 + 3]));
        __syncwarp();
        {
          const int col4 = lane & 3, rsub = lane >> 2;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int row = i * 8 + rsub;
            if (n_base + row < N)
              *reinterpret_cast<float4*>(part + (size_t)Npad2 * KP + (size_t)(n_base + row) * 32 + half * 16 + col4 * 4) =
                  *reinterpret_cast<const float4*>(scr + row * PITCH + col4 * 4);
          }
        }
        __syncwarp();
      }
    }
    fence_before();
  }
  if (warp == 0) DP_PHASE(2, 61);
  __syncthreads();
  if (warp == 0) DP_PHASE(2, 62);
  if (warp == 12) tmem_dealloc(tmem_base, 512);
}

// dP[n][k] = sum over CTAs; dW[c][n] = sum over CTAs (stored [n][32] per CTA).  With fin.scalars set this is
// also the step's epilogue: dW += w_l1 * sign(W) off the class identity, block 0 writes the eight scalars.
__global__ void __launch_bounds__(256)
k_reduce_partials(const float* __restrict__ partials, int n_parts, int Npad2, int KP, int N, int K,
                  int C, float* __restrict__ dP, float* __restrict__ dW, const float* __restrict__ W,
                  const int32_t* __restrict__ proto_class, const StepFinalize fin) {
  // the next step's prologue kernel may start its label histogram now (it waits for this grid before it writes)
  asm volatile("griddepcontrol.launch_dependents;");
  // 4 threads per output element, each sums a quarter of the per-CTA partials
  const size_t per = (size_t)Npad2 * KP + (size_t)Npad2 * 32;
  const int total = N * K + C * N;
  const int idx = (blockIdx.x * 256 + threadIdx.x) >> 2, q4 = threadIdx.x & 3;
  float a = 0.f, extra = 0.f;
  float* dst = nullptr;
  if (idx < total) {
    size_t off;
    if (idx < N * K) {
      const int n = idx / K, k = idx - n * K;
      off = (size_t)n * KP + k;
      dst = dP ? dP + idx : nullptr;
    } else {
      const int j = idx - N * K, c = j / N, n = j - c * N;
      off = (size_t)Npad2 * KP + (size_t)n * 32 + c;
      dst = dW ? dW + j : nullptr;
      if (fin.scalars != nullptr && q4 == 0 && dst != nullptr && proto_class[n] != c) {
        const float w = W[j];
        extra = fin.w_l1 * fin.grad_scale * ((w > 0.f) ? 1.f : ((w < 0.f) ? -1.f : 0.f));
      }
    }
    if (dst != nullptr) {
      // eight partials in flight per thread (the 37 loads of a thread were two per round trip)
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
      const float* src = partials + off;
      int q = q4;
      for (; q + 28 < n_parts; q += 32) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = src[(size_t)(q + 4 * u) * per];
        a0 += v[0] + v[4]; a1 += v[1] + v[5]; a2 += v[2] + v[6]; a3 += v[3] + v[7];
      }
      {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = (q + 4 * u < n_parts) ? src[(size_t)(q + 4 * u) * per] : 0.f;
        a0 += v[0] + v[4]; a1 += v[1] + v[5]; a2 += v[2] + v[6]; a3 += v[3] + v[7];
      }
      a = (a0 + a1) + (a2 + a3);
    }
  }
  a += __shfl_xor_sync(0xffffffffu, a, 1);
  a += __shfl_xor_sync(0xffffffffu, a, 2);

  // ---- data-parallel: the all-reduce of [dP | dW | loss sums] happens here, over peer-mapped memory
  const DpComm& cm = fin.comm;
  const DpLayout& lay = fin.lay;
  __shared__ int s_timeout;
  const int par = (int)(cm.seq & 1u), Wd = cm.world;
  if (cm.seq != 0u) {
    if (threadIdx.x == 0) s_timeout = 0;
    if (idx < total) {
      for (int r = q4; r < Wd; r += 4) {
        volatile float* slot = reinterpret_cast<volatile float*>(cm.peer[r] + lay.grad) +
                               (size_t)(par * Wd + cm.rank) * lay.n_grad_pad;
        slot[idx] = a;
      }
    }
    if (blockIdx.x == 0 && threadIdx.x < 4 * Wd) {
      const int r = threadIdx.x >> 2, k = threadIdx.x & 3;
      volatile double* slot = reinterpret_cast<volatile double*>(cm.peer[r] + lay.acc) + (size_t)(par * Wd + cm.rank) * 4;
      slot[k] = fin.accum[k];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < Wd) {
      const int r = threadIdx.x;
      st_release_sys(reinterpret_cast<unsigned int*>(cm.peer[r] + lay.gflag) +
                         (size_t)(par * Wd + cm.rank) * lay.n_blocks + blockIdx.x, cm.seq);
      const bool ok = dp_wait_flag(reinterpret_cast<const unsigned int*>(cm.peer[cm.rank] + lay.gflag) +
                                       (size_t)(par * Wd + r) * lay.n_blocks + blockIdx.x, cm.seq, cm.timeout_ns);
      if (!ok) s_timeout = 1;
    }
    __syncthreads();
    if (idx < total && q4 == 0) {
      const volatile float* mine = reinterpret_cast<const volatile float*>(cm.peer[cm.rank] + lay.grad) +
                                   (size_t)par * Wd * lay.n_grad_pad + idx;
      float t = 0.f;
      for (int r = 0; r < Wd; ++r) t += mine[(size_t)r * lay.n_grad_pad];
      a = s_timeout ? __int_as_float(0x7fc00000) : t;
    }
  }
  if (q4 == 0 && dst != nullptr) *dst = a + extra;

  if (fin.scalars != nullptr && blockIdx.x == 0) {
    __shared__ float s_part[8];
    float l1 = 0.f;
    for (int i0 = 0; i0 < C * N; i0 += 256 * 8) {          // both loads of eight elements in flight at once
      float wv[8];
      int pv[8], cv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * 256 + (int)threadIdx.x;
        const bool in = i < C * N;
        const int c = in ? i / N : 0, n = in ? i - c * N : 0;
        cv[u] = c;
        pv[u] = in ? proto_class[n] : c;
        wv[u] = in ? W[i] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (pv[u] != cv[u]) l1 += fabsf(wv[u]);
    }
    l1 = warp_sum(l1);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = l1;
    __syncthreads();
    if (threadIdx.x == 0) {
      double tot = 0.0;
      for (int w = 0; w < 8; ++w) tot += (double)s_part[w];
      double acc[4] = {fin.accum[0], fin.accum[1], fin.accum[2], fin.accum[3]};
      if (cm.seq != 0u) {
        const volatile double* mine = reinterpret_cast<const volatile double*>(cm.peer[cm.rank] + lay.acc) +
                                      (size_t)par * Wd * 4;
        for (int k = 0; k < 4; ++k) {
          double t = 0.0;
          for (int r = 0; r < Wd; ++r) t += mine[r * 4 + k];
          acc[k] = t;
        }
        if (s_timeout) acc[3] += 1.0;
      }
      const double n_valid = (double)fin.counts[0], n_terms = (double)fin.counts[1];
      const double ce = n_valid > 0 ? acc[0] / n_valid
                                     : __longlong_as_double(0x7ff8000000000000LL);  // empty mean = NaN, as torch
      const double kld = n_terms > 0 ? acc[2] / n_terms : 0.0;
      fin.scalars[0] = (double)fin.w_ce * ce + (double)fin.w_kld * kld + (double)fin.w_l1 * tot;
      fin.scalars[1] = ce;
      fin.scalars[2] = kld;
      fin.scalars[3] = tot;
      fin.scalars[4] = acc[1];
      fin.scalars[5] = n_valid;
      fin.scalars[6] = n_terms;
      fin.scalars[7] = acc[3];
    }
  }
}

#include "tc_bwd_fused.cuh"

int g_sms = 0;
int sm_count() {
  if (g_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_sms <= 0) g_sms = 148;
  }
  return g_sms;
}

bool tc_bwd_supported(const HeadShape& s, int prec_mode) {
  return prec_mode == ADLM_PREC_TF32 && s.K <= 64 && s.N <= 192 && s.C <= 32;
}

// The fused backward (tc_bwd_fused.cuh) is selected by ADLM_BWD_FUSED=1 / 0 in the environment; the default below is
// whichever of the two measured faster on the B200 (profiles/r2_*).  ADLM_BWD_TWO_KERNELS=1 also forces the two-kernel form.
constexpr bool FUSED_BWD_DEFAULT = false;
bool use_fused_bwd() {
  static const bool v = [] {
    const char* two = getenv("ADLM_BWD_TWO_KERNELS");
    if (two != nullptr && two[0] == '1') return false;
    const char* e = getenv("ADLM_BWD_FUSED");
    if (e != nullptr && (e[0] == '0' || e[0] == '1')) return e[0] == '1';
    return FUSED_BWD_DEFAULT;
  }();
  return v;
}

template <int KP, int NB>
int launch_fused(TcBwdParams prm, const StepFinalize& fin, float* dP, float* dW, cudaStream_t st) {
  using L = FuSmem<KP, NB>;
  const size_t bytes = L::total(prm.C8);
  if (bytes > 227 * 1024) {
    set_error("tc_head_bwd: %zu bytes of shared memory needed (> 227 KB) for K=%d N=%d C=%d", bytes, prm.K, prm.N, prm.C);
    return ADLM_E_UNSUPPORTED;
  }
  prm.num_tiles = (int)((prm.M + FT - 1) / FT);
  {
    static const int flags = [] { const char* e = getenv("ADLM_FUSED_EXP"); return e ? atoi(e) : 0; }();
    prm.exp_flags = flags;
  }
  const int grid = prm.num_tiles < sm_count() ? prm.num_tiles : sm_count();
  const bool after_kld = prm.own_grad != nullptr;      // step path: the launch before this one is k_kld
  ADLM_CUDA(cudaFuncSetAttribute(k_tc_bwd_fused<KP, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  ADLM_CUDA(launch_maybe_pdl(after_kld, k_tc_bwd_fused<KP, NB>, dim3(grid), dim3(FU_THREADS), bytes, st, prm));
  ADLM_LAUNCH_CHECK();
  const int total = prm.N * prm.K + prm.C * prm.N;
  k_reduce_partials<<<(total * 4 + 255) / 256, 256, 0, st>>>(prm.partials, grid, NB * 128, KP, prm.N, prm.K, prm.C, dP, dW,
                                                             prm.W, prm.proto_class, fin);
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

template <int KP>
int launch_bwd(TcBwdParams prm, float* dP, float* dW, const StepFinalize& fin, void* workspace, size_t workspace_bytes,
               cudaStream_t st) {
  const int grid = prm.num_tiles < sm_count() ? prm.num_tiles : sm_count();
  KernelTimer timer(ADLM_K_HEAD_BWD, st);
  // dZ together with dP / dW and no upstream gradient on the distance / similarity maps (the training step, and
  // autograd through the logits): the fused kernel forms G once for all three
  if (use_fused_bwd() && prm.dz != nullptr && (dP != nullptr || dW != nullptr) && prm.dsim == nullptr &&
      prm.ddist == nullptr && prm.C8 <= 24) {
    const int NB = (prm.N + 127) / 128, Npad2 = NB * 128;
    const size_t per = ((size_t)Npad2 * KP + (size_t)Npad2 * 32) * sizeof(float);
    if (workspace == nullptr || workspace_bytes < per * grid) {
      set_error("tc_head_bwd: workspace too small (%zu < %zu)", workspace_bytes, per * grid);
      return ADLM_E_WORKSPACE;
    }
    prm.partials = static_cast<float*>(workspace);
    return NB == 1 ? launch_fused<KP, 1>(prm, fin, dP, dW, st) : launch_fused<KP, 2>(prm, fin, dP, dW, st);
  }
  bool dz_launched = false;
  if (prm.dz != nullptr) {
    using L = DzSmem<KP>;
    const size_t bytes = L::total(prm.Npad, prm.C8);
    const bool after_kld = prm.own_grad != nullptr;      // step path: the launch before this one is k_kld
    if (prm.dsim != nullptr || prm.ddist != nullptr) {
      ADLM_CUDA(cudaFuncSetAttribute(k_tc_bwd_dz<KP, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
      ADLM_CUDA(launch_maybe_pdl(after_kld, k_tc_bwd_dz<KP, true>, dim3(grid), dim3(256), bytes, st, prm));
    } else {
      ADLM_CUDA(cudaFuncSetAttribute(k_tc_bwd_dz<KP, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
      ADLM_CUDA(launch_maybe_pdl(after_kld, k_tc_bwd_dz<KP, false>, dim3(grid), dim3(256), bytes, st, prm));
    }
    ADLM_LAUNCH_CHECK();
    dz_launched = true;
  }
  if (dP != nullptr || dW != nullptr) {
    using L = DpSmem<KP>;
    const int NB = (prm.N + 127) / 128, Npad2 = NB * 128;
    const size_t per = ((size_t)Npad2 * KP + (size_t)Npad2 * 32) * sizeof(float);
    if (workspace == nullptr || workspace_bytes < per * grid) {
      set_error("tc_head_bwd: workspace too small (%zu < %zu)", workspace_bytes, per * grid);
      return ADLM_E_WORKSPACE;
    }
    prm.partials = static_cast<float*>(workspace);
    const size_t bytes = L::total(Npad2, prm.C8);
    const bool generic = prm.dsim != nullptr || prm.ddist != nullptr;
#define ADLM_DP_LAUNCH(NBV, GEN)                                                                                       \
  do {                                                                                                                 \
    ADLM_CUDA(cudaFuncSetAttribute(k_tc_bwd_dp<KP, NBV, GEN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)); \
    ADLM_CUDA(launch_maybe_pdl(dz_launched, k_tc_bwd_dp<KP, NBV, GEN>, dim3(grid), dim3(DP_THREADS), bytes, st, prm));   \
  } while (0)
    if (NB == 1) { if (generic) ADLM_DP_LAUNCH(1, true); else ADLM_DP_LAUNCH(1, false); }
    else         { if (generic) ADLM_DP_LAUNCH(2, true); else ADLM_DP_LAUNCH(2, false); }
#undef ADLM_DP_LAUNCH
    ADLM_LAUNCH_CHECK();
    const int total = prm.N * prm.K + prm.C * prm.N;
    k_reduce_partials<<<(total * 4 + 255) / 256, 256, 0, st>>>(prm.partials, grid, Npad2, KP, prm.N, prm.K, prm.C, dP, dW,
                                                               prm.W, prm.proto_class, fin);
    ADLM_LAUNCH_CHECK();
  }
  return ADLM_OK;
}

}  // namespace

#ifdef ADLM_PHASE_CLOCKS
extern "C" ADLM_API int adlm_debug_dz_phase_clocks(long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, g_dz_phase, sizeof(long long) * 128) == cudaSuccess ? 0 : -3;
}
extern "C" ADLM_API int adlm_debug_dp_phase_clocks(long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, g_dp_phase, sizeof(long long) * 192) == cudaSuccess ? 0 : -3;
}
#endif

size_t tc_bwd_workspace_bytes(int K, int N) {
  const int NB = (N + 127) / 128, Npad2 = NB * 128, KP = K <= 32 ? 32 : 64;
  return ((size_t)Npad2 * KP + (size_t)Npad2 * 32) * sizeof(float) * (size_t)sm_count();
}

int tc_head_bwd(const HeadShape& s, int prec_mode, const float* z, const float* P, const float* W,
                const float* dlogits, const float* ddist, const float* dsim, float* dz, float* dP,
                float* dW, const StepExtras& ex, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  if (!tc_bwd_supported(s, prec_mode)) {
    if (prec_mode == ADLM_PREC_TF32) {
      // beyond the tensor-core backward's operand budget: refused like the forward does, the caller picks
      // ADLM_PREC_FP32 explicitly — no silent change of kernel class
      set_error("adlm_head_bwd: the tcgen05 backward needs K <= 64, N <= 192, C <= 32 (got K=%d N=%d C=%d); use ADLM_PREC_FP32",
                s.K, s.N, s.C);
      return ADLM_E_UNSUPPORTED;
    }
    // ADLM_PREC_TF32X3 is the fp32-class mode: its backward IS the fp32 CUDA-core backward, for every shape
    // (include/adlm_head.h) — same mathematics, tighter arithmetic
    if (ex.dz_sigmoid) {
      set_error("tc_head_bwd: the fused sigmoid backward needs ADLM_PREC_TF32, K <= 64, N <= 192, C <= 32");
      return ADLM_E_UNSUPPORTED;
    }
    return simt_head_bwd(s, z, P, W, dlogits, ddist, dsim, dz, dP, dW, ex, st);
  }
  TcBwdParams prm;
  prm.z = z; prm.P = P; prm.W = W;
  prm.dlogits = dlogits; prm.ddist = ddist; prm.dsim = dsim;
  prm.target = ex.own_grad ? ex.target : nullptr;
  prm.proto_class = ex.proto_class; prm.proto_rank = ex.proto_rank; prm.own_grad = ex.own_grad;
  prm.dz = dz; prm.partials = nullptr;
  prm.dz_sigmoid = ex.dz_sigmoid;
  prm.B = s.B; prm.K = s.K; prm.HW = s.HW; prm.N = s.N; prm.C = s.C;
  prm.eps = s.eps; prm.act_mode = s.act_mode;
  prm.Npad = (s.N + 31) & ~31;
  prm.C8 = (s.C + 7) & ~7;
  prm.M = (long long)s.B * s.HW;
  prm.num_tiles = (int)((prm.M + TILE - 1) / TILE);
  prm.exp_flags = 0;
  if (s.K <= 32) return launch_bwd<32>(prm, dP, dW, ex.fin, workspace, workspace_bytes, st);
  return launch_bwd<64>(prm, dP, dW, ex.fin, workspace, workspace_bytes, st);
}

}  // namespace adlm
