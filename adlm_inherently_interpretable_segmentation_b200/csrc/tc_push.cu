// Push search on the tensor cores (ADLM_PREC_TF32 / ADLM_PREC_TF32X3): segmentation/push.py:249-280 as one
// streaming, warp-specialised kernel.
//
// Transposed problem: TMEM lanes = prototypes, columns = the 128 patches of a tile,
//   D^T[n][px] = Pc[n] . Zc[px]      (3xTF32 = hi*lo + lo*hi + hi*hi: fp32-class distances, so the argmin matches)
// The A operand (prototypes, constant for the whole launch) is written ONCE into TMEM and every MMA is of the
// .ts form (A from TMEM, B = feature tile from shared memory): half the shared-memory operand traffic of an
// SS MMA and no prototype copy in shared memory, which leaves room for three feature-tile stages.
// Every epilogue thread owns ONE prototype and keeps its running (min distance, patch id) in two registers for
// all tiles of the CTA: no shuffles, no shared-memory reduction, one 64-bit atomicMin per prototype per CTA at
// the very end.  Patches are visited in increasing global id, so a strict '<' keeps the earliest patch
// (push.py:265); the class filter is one AND against the patch's class bitmask (push.py:216-223).
//
// Roles (512 threads): warps 0-6 epilogue (TMEM lane quarter = warp % 4, prototype block = warp / 4; N <= 224),
// warp 7 MMA issuer, warps 8-15 producer (two threads per patch, half of the K features each: LDG two tiles
// ahead into a double register buffer, centre, round / split, write the B operand, partial |z|^2, mask).
// TMEM: [A hi | A lo] per prototype block (4 x 64 columns), one 128-column accumulator per block: the epilogue
// of block 0 overlaps the MMAs of block 1 and vice versa.
#include "tc_head.cuh"
#include "tc_common.cuh"

namespace adlm {

namespace {

using namespace tc;

constexpr int TILE = 128;
constexpr float CENTRE = 0.5f;
constexpr int PUSH_THREADS = 512;

struct TcPushParams {
  const float* z;
  const uint32_t* mask;
  const float* P;
  const int32_t* proto_class;
  unsigned long long* packed_min;
  unsigned long long patch_base;
  int K, HW, N;
  int M;          // n_img * HW
  int num_tiles;
  // per-image minima (adlm_push_scan_images): tiles never straddle images (tiles_per_img each, the last one
  // partly empty), every CTA owns a contiguous tile range so that an image change is rare, packed_min is
  // [n_img][N] and the stored id is the patch index inside its image
  int per_image;
  int tiles_per_img;
};

// tile `it` of this CTA -> first global patch m0 and the number of real patches in the tile
struct TileMap {
  int first, stride, tpi, HW, M, per_image;
  __device__ __forceinline__ void get(int it, int& m0, int& valid) const {
    const int t = first + it * stride;
    if (per_image) {
      const int img = t / tpi, off = (t - img * tpi) * TILE;
      m0 = img * HW + off;
      valid = min(TILE, HW - off);
    } else {
      m0 = t * TILE;
      valid = min(TILE, M - m0);
    }
  }
};

template <int KP, bool X3>
struct PushSmem {
  static constexpr int PARTS = X3 ? 2 : 1;
  static constexpr int STAGES = (X3 && KP > 32) ? 2 : 3;
  // operand tile (stage s, part): [KP/4][128 px][4]
  static constexpr __host__ __device__ size_t zb(int stage, int part) { return ((size_t)stage * PARTS + part) * KP * TILE * 4; }
  static constexpr __host__ __device__ size_t misc() { return zb(STAGES, 0); }
  // class tables [2 slots][33 rows][132] f32, x2 halves [2][128] f32, 16 barriers, tmem base
  static constexpr __host__ __device__ size_t total() { return misc() + 2 * 33 * 132 * 4 + 1024 + 128 + 16; }
};

// barrier indices
constexpr int B_FULL = 0;        // [3] operand stage written (256 producer threads)
constexpr int B_SFREE = 3;       // [3] MMAs reading the stage are done (commit)
constexpr int B_AFULL = 6;       // [2] accumulator of block ready (commit)
constexpr int B_AEMPTY = 8;      // [2] accumulator of block drained (epilogue warps of the block)
constexpr int B_META = 10;       // [2] |z|^2 / mask slot written (256 producer threads)
constexpr int B_MEMPTY = 12;     // [2] |z|^2 / mask slot drained (the 128 or 224 epilogue threads that own a block)

template <int KP, bool X3, int NB>
__global__ void __launch_bounds__(PUSH_THREADS, 1) k_tc_push(const TcPushParams p) {
  using L = PushSmem<KP, X3>;
  constexpr int STAGES = L::STAGES;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int K = p.K, N = p.N, HW = p.HW;
  uint8_t* misc = smem + L::misc();
  // xsc[slot][c][px] = |z_px|^2 if class c occurs in patch px, else +inf (row 32: all +inf, for prototypes
  // without a valid class).  One FFMA per (prototype, patch) then yields a distance that can only win if the
  // patch is a candidate of the prototype's class.
  constexpr int XROW = 132, XSLOT = 33 * XROW;
  float* xsc = reinterpret_cast<float*>(misc);                                  // [2][33][132]
  float* x2h = reinterpret_cast<float*>(misc + 2 * XSLOT * 4);                  // [2 halves][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(misc + 2 * XSLOT * 4 + 1024);
  uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(bars + 16);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (warp == 7) tmem_alloc(smem_u32(tmem_base_s), 512);
  if (tid == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(smem_u32(&bars[B_FULL + i]), 256);
      mbar_init(smem_u32(&bars[B_SFREE + i]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&bars[B_AFULL + i]), 1);
      mbar_init(smem_u32(&bars[B_AEMPTY + i]), i == 0 ? 128 : 96);
      mbar_init(smem_u32(&bars[B_META + i]), 256);
      mbar_init(smem_u32(&bars[B_MEMPTY + i]), NB == 2 ? 224 : 128);   // the epilogue threads that have a block
    }
    mbar_init_fence();
  }
  const uint32_t bar0 = smem_u32(&bars[0]);
  auto bar = [&](int idx) { return bar0 + (uint32_t)idx * 8u; };
  TileMap tmap;
  tmap.tpi = p.tiles_per_img; tmap.HW = HW; tmap.M = p.M; tmap.per_image = p.per_image;
  int iters;
  if (p.per_image) {
    const long long lo = (long long)blockIdx.x * p.num_tiles / (int)gridDim.x;
    const long long hi = (long long)(blockIdx.x + 1) * p.num_tiles / (int)gridDim.x;
    tmap.first = (int)lo; tmap.stride = 1; iters = (int)(hi - lo);
  } else {
    tmap.first = (int)blockIdx.x; tmap.stride = (int)gridDim.x;
    iters = p.num_tiles > (int)blockIdx.x ? (p.num_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  }

  // producer warps start fetching their first two tiles right away
  const bool is_producer = warp >= 8;
  const int cw = (tid - 256) & 127, half = (tid - 256) >> 7;     // producer: patch column, feature half
  constexpr int KH = KP / 2;
  float zrA[KH], zrB[KH];
  uint32_t mkA = 0u, mkB = 0u;
  auto prefetch = [&](int it, float (&zr)[KH], uint32_t& mk) {
    int m0 = 0, valid = 0;
    if (it < iters) tmap.get(it, m0, valid);
    const int m = m0 + cw;
    if (cw < valid) {
      const int img = m / HW;
      const float* src = p.z + ((size_t)img * K + half * KH) * HW + (m - img * HW);
      if (K == KP) {
#pragma unroll
        for (int k = 0; k < KH; ++k) { zr[k] = __ldg(src); src += HW; }
      } else {
#pragma unroll
        for (int k = 0; k < KH; ++k) zr[k] = (half * KH + k < K) ? __ldg(src + (size_t)k * HW) : CENTRE;
      }
      mk = p.mask ? __ldg(p.mask + m) : 0xffffffffu;       // no mask: every patch is a candidate
    } else {
#pragma unroll
      for (int k = 0; k < KH; ++k) zr[k] = CENTRE;
      mk = 0u;
    }
  };
  if (is_producer) {
    prefetch(0, zrA, mkA);
    prefetch(1, zrB, mkB);
  }
  fence_before();
  __syncthreads();                 // TMEM base address, barriers
  fence_after();
  const uint32_t tmem_base = *tmem_base_s;
  // TMEM columns: A operand of block b at b*128 (hi) and b*128 + 64 (lo); accumulator of block b at 256 + b*128
  const uint32_t tA = tmem_base, tD = tmem_base + 256;

  // ---- epilogue threads write their prototype (centred, rounded / split) into the TMEM A operand ----
  float p2 = 0.f;
  int crow = 32;                   // class-table row of this prototype
  const int q = warp & 3, blk = warp >> 2;
  const int n = blk * 128 + q * 32 + lane;
  const uint32_t lane_field = (uint32_t)(q * 32) << 16;
  const bool is_epilogue = warp < 7;
  if (is_epilogue && blk < NB) {
    const float* row = p.P + (size_t)(n < N ? n : 0) * K;
#pragma unroll
    for (int c = 0; c < KP / 16; ++c) {
      uint32_t hi[16], lo[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int k = c * 16 + i;
        const float v = (n < N && k < K) ? __ldg(row + k) - CENTRE : 0.f;
        const float h = rna_tf32_mask(v);
        const float w = X3 ? v : h;
        p2 = fmaf(w, w, p2);
        hi[i] = __float_as_uint(h);
        lo[i] = __float_as_uint(X3 ? rna_tf32_mask(v - h) : 0.f);
      }
      tmem_st16(tA + lane_field + blk * 128 + c * 16, hi);
      if (X3) tmem_st16(tA + lane_field + blk * 128 + 64 + c * 16, lo);
    }
    tmem_st_wait();
    if (n < N) {
      const int pc = p.proto_class ? p.proto_class[n] : 0;   // no class table: every prototype reads bit 0
      if (pc >= 0 && pc < 32) crow = pc;
    }
  }
  for (int i = tid; i < 2 * XROW; i += PUSH_THREADS)
    xsc[(i / XROW) * XSLOT + 32 * XROW + (i % XROW)] = __int_as_float(0x7f800000);
  fence_before();
  __syncthreads();
  fence_after();

  if (is_producer) {
    // ================= producer: thread = (patch, feature half) =================
    auto produce = [&](int it, const float (&zr)[KH], uint32_t mk) {
      const int a = it & 1, j = it >> 1;
      const int sg = it % STAGES, u = it / STAGES;
      float* ZB_hi = reinterpret_cast<float*>(smem + L::zb(sg, 0));
      float* ZB_lo = reinterpret_cast<float*>(smem + L::zb(sg, X3 ? 1 : 0));
      if (u > 0) mbar_wait(bar(B_SFREE + sg), (uint32_t)(u - 1) & 1u);        // operand stage consumed
      if (j > 0) mbar_wait(bar(B_MEMPTY + a), (uint32_t)(j - 1) & 1u);        // |z|^2 / mask slot drained
      float x2 = 0.f;
#pragma unroll
      for (int k4 = 0; k4 < KH / 4; ++k4) {
        float hi[4], lo[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float v = zr[k4 * 4 + e] - CENTRE;
          hi[e] = rna_tf32_mask(v);
          lo[e] = X3 ? rna_tf32_mask(v - hi[e]) : 0.f;
          const float w = X3 ? v : hi[e];
          x2 = fmaf(w, w, x2);
        }
        const size_t off = ((size_t)(half * (KH / 4) + k4) * TILE + cw) * 4;
        *reinterpret_cast<float4*>(ZB_hi + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
        if (X3) *reinterpret_cast<float4*>(ZB_lo + off) = make_float4(lo[0], lo[1], lo[2], lo[3]);
      }
      fence_async_smem();
      mbar_arrive(bar(B_FULL + sg));
      // |z|^2 = sum of the two halves' partial sums, then this thread fills 16 of the 32 class rows
      x2h[half * 128 + cw] = x2;
      named_bar_sync(1, 256);
      const float x2t = x2h[cw] + x2h[128 + cw];
      named_bar_sync(1, 256);                      // x2h is rewritten by the next tile
      float* tab = xsc + a * XSLOT + (half * 16) * XROW + cw;
      const uint32_t mh = mk >> (half * 16);
#pragma unroll
      for (int c = 0; c < 16; ++c) tab[c * XROW] = ((mh >> c) & 1u) ? x2t : __int_as_float(0x7f800000);
      mbar_arrive(bar(B_META + a));
    };
    for (int it = 0; it < iters; it += 2) {
      produce(it, zrA, mkA);
      prefetch(it + 2, zrA, mkA);
      if (it + 1 < iters) {
        produce(it + 1, zrB, mkB);
        prefetch(it + 3, zrB, mkB);
      }
    }
  } else if (warp == 7) {
    // ================= MMA issuer =================
    if (lane == 0) {
      const uint32_t idesc = make_idesc_tf32(128, TILE);
      const int ksteps = (K + 7) >> 3;
      const uint32_t b_kstep = (2u * TILE * 16u) >> 4;
      for (int it = 0; it < iters; ++it) {
        const int sg = it % STAGES, u = it / STAGES;
        const uint64_t db_hi = make_desc(smem_u32(smem + L::zb(sg, 0)), TILE * 16, 128);
        const uint64_t db_lo = make_desc(smem_u32(smem + L::zb(sg, X3 ? 1 : 0)), TILE * 16, 128);
        mbar_wait(bar(B_FULL + sg), (uint32_t)u & 1u);
#pragma unroll
        for (int b = 0; b < NB; ++b) {
          if (it > 0) mbar_wait(bar(B_AEMPTY + b), (uint32_t)(it - 1) & 1u);  // previous tile's epilogue of this block
          fence_after();
          const uint32_t d = tD + b * 128, a_hi = tA + b * 128, a_lo = a_hi + 64;
#pragma unroll
          for (int ks = 0; ks < KP / 8; ++ks) {
            if (ks < ksteps) {
              const uint64_t bo = (uint64_t)(ks * b_kstep);
              if (X3) {
                mma_ts(d, a_hi + ks * 8, db_lo + bo, idesc, ks > 0 ? 1u : 0u);
                mma_ts(d, a_lo + ks * 8, db_hi + bo, idesc, 1u);
                mma_ts(d, a_hi + ks * 8, db_hi + bo, idesc, 1u);
              } else {
                mma_ts(d, a_hi + ks * 8, db_hi + bo, idesc, ks > 0 ? 1u : 0u);
              }
            }
          }
          mma_commit(bar(B_AFULL + b));
        }
        mma_commit(bar(B_SFREE + sg));
      }
    }
  } else {
    // ================= epilogue: thread = prototype =================
    const bool warp_has_work = blk < NB && (blk * 128 + q * 32) < N;   // warps whose prototypes are all padding skip the math
    float best = __int_as_float(0x7f800000);                            // +inf
    int best_m = -1;
    const float negp2 = -p2;
    int cur_img = -1;
    auto flush = [&]() {
      if (n < N && best_m >= 0) {
        const unsigned long long dbits = (unsigned long long)__float_as_uint(fmaxf(best + p2, 0.f)) << 32;
        if (p.per_image)
          atomicMin(p.packed_min + (size_t)cur_img * N + n, dbits | (unsigned long long)(best_m - cur_img * HW));
        else
          atomicMin(p.packed_min + n, dbits | (p.patch_base + (unsigned long long)best_m));
      }
    };
    for (int it = 0; it < iters; ++it) {
      const int a = it & 1, j = it >> 1;
      int m0, valid;
      tmap.get(it, m0, valid);
      if (p.per_image) {
        const int img = m0 / HW;
        if (img != cur_img) {
          if (cur_img >= 0) flush();
          best = __int_as_float(0x7f800000);
          best_m = -1;
          cur_img = img;
        }
      }
      // Warps without a prototype block (NB == 1: warps 4-6) take no part in the slot hand-shake.  They used to
      // arrive on B_MEMPTY without waiting for anything, ran ahead, and their arrivals for a LATER tile completed
      // the current phase early: with N <= 128 and three or more tiles per CTA the producer then overwrote a
      // |z|^2 / class table that warp 0 was still reading (wrong minima; caught by the per-image scan test).
      if (blk < NB) {
        mbar_wait(bar(B_META + a), (uint32_t)j & 1u);
        mbar_wait(bar(B_AFULL + blk), (uint32_t)it & 1u);
        fence_after();
        if (warp_has_work) {
          const uint32_t t0 = tD + blk * 128 + lane_field;
          const float* xrow = xsc + a * XSLOT + crow * XROW;
          uint32_t ra[16], rb[16];
          // t = |z|^2 - 2 z.p (or +inf for non-candidates); d = max(t + |p|^2, 0) is monotone in t, so the search
          // runs on tc = max(t, -|p|^2) and |p|^2 is added once per prototype at the end
          auto chunk = [&](int c0, const uint32_t (&r)[16]) {
#pragma unroll
            for (int v4 = 0; v4 < 4; ++v4) {
              const float4 xq = *reinterpret_cast<const float4*>(xrow + c0 + v4 * 4);
              const float xs[4] = {xq.x, xq.y, xq.z, xq.w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int i = v4 * 4 + e;
                const float tc = fmaxf(fmaf(-2.f, __uint_as_float(r[i]), xs[e]), negp2);
                if (tc < best) best_m = m0 + c0 + i;
                best = fminf(best, tc);
              }
            }
          };
          tmem_ld16(t0, ra);
          for (int c0 = 0; c0 < TILE; c0 += 32) {
            tmem_ld_wait();
            tmem_ld16(t0 + c0 + 16, rb);
            chunk(c0, ra);
            tmem_ld_wait();
            if (c0 + 32 < TILE) tmem_ld16(t0 + c0 + 32, ra);
            chunk(c0 + 16, rb);
          }
        }
        fence_before();
        mbar_arrive(bar(B_AEMPTY + blk));
        mbar_arrive(bar(B_MEMPTY + a));
      }
    }
    // padding columns of a partly empty tile carry mask 0 and never win
    flush();
  }
  fence_before();
  __syncthreads();
  if (warp == 7) tmem_dealloc(tmem_base, 512);
}

int g_sms_push = 0;
int sm_count_push() {
  if (g_sms_push == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sms_push, cudaDevAttrMultiProcessorCount, dev);
    if (g_sms_push <= 0) g_sms_push = 148;
  }
  return g_sms_push;
}

template <int KP, bool X3, int NB>
int launch_push(const TcPushParams& prm, cudaStream_t st) {
  using L = PushSmem<KP, X3>;
  const int grid = prm.num_tiles < sm_count_push() ? prm.num_tiles : sm_count_push();
  const size_t bytes = L::total();
  static_assert(L::total() <= 227 * 1024, "push kernel shared memory");
  ADLM_CUDA(cudaFuncSetAttribute(k_tc_push<KP, X3, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  KernelTimer timer(ADLM_K_PUSH_SCAN, st);
  k_tc_push<KP, X3, NB><<<grid, PUSH_THREADS, bytes, st>>>(prm);
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

}  // namespace

int tc_push_scan(const float* z, int n_img, int K, int HW, const uint32_t* mask, const float* P, int N,
                 const int32_t* proto_class, long long patch_base, int prec_mode,
                 unsigned long long* packed_min, int per_image, cudaStream_t st) {
  const long long M = (long long)n_img * HW;
  const int tiles_per_img = (HW + TILE - 1) / TILE;
  // Shapes beyond the tensor-core kernel's operand budget are refused, like the forward does: the caller picks
  // ADLM_PREC_FP32 (the CUDA-core scan: same search, same keys) explicitly — no silent change of kernel class.
  if (K > 64 || N > 224 || M >= (1ll << 31) - 256 || (long long)n_img * tiles_per_img >= (1ll << 31)) {
    set_error("adlm_push_scan: the tcgen05 scan needs K <= 64 and N <= 224 (got K=%d N=%d); use ADLM_PREC_FP32", K, N);
    return ADLM_E_UNSUPPORTED;
  }
  TcPushParams prm;
  prm.z = z; prm.mask = mask; prm.P = P; prm.proto_class = proto_class; prm.packed_min = packed_min;
  prm.patch_base = (unsigned long long)patch_base;
  prm.K = K; prm.HW = HW; prm.N = N;
  prm.M = (int)M;
  prm.per_image = per_image;
  prm.tiles_per_img = tiles_per_img;
  prm.num_tiles = per_image ? n_img * tiles_per_img : (int)((M + TILE - 1) / TILE);
  const bool x3 = prec_mode == ADLM_PREC_TF32X3;
  const int nb = N > 128 ? 2 : 1;
#define ADLM_PUSH_CASE(KPV)                                                                    \
  do {                                                                                         \
    if (x3) return nb == 2 ? launch_push<KPV, true, 2>(prm, st) : launch_push<KPV, true, 1>(prm, st);   \
    return nb == 2 ? launch_push<KPV, false, 2>(prm, st) : launch_push<KPV, false, 1>(prm, st);         \
  } while (0)
  if (K <= 32) ADLM_PUSH_CASE(32);
  ADLM_PUSH_CASE(64);
#undef ADLM_PUSH_CASE
}

}  // namespace adlm
