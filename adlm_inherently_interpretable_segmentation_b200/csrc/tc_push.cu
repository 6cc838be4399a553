// Push search on the tensor cores (ADLM_PREC_TF32 / ADLM_PREC_TF32X3): segmentation/push.py:249-280 as one
// streaming, warp-specialised kernel.
//
// Transposed problem: TMEM lanes = prototypes, columns = the 128 patches of a tile,
//   D^T[n][px] = Pc[n] . Zc[px]      (A = prototypes, staged once per CTA; B = the feature tile; 3xTF32 =
//                                     hi*lo + lo*hi + hi*hi, fp32-class distances so that the argmin matches)
// so every epilogue thread owns ONE prototype and keeps its running (min distance, patch id) in two registers
// for all tiles of the CTA: no shuffles, no shared-memory reduction, one 64-bit atomicMin per prototype per CTA
// at the very end.  Patches are visited in increasing global id, so a strict '<' keeps the earliest patch
// (push.py:265); the class filter is one AND against the patch's class bitmask (push.py:216-223).
//
// Roles (416 threads): warps 0-7 epilogue (lane quarter = warp % 4, prototype block / column half = warp / 4),
// warps 8-11 producer (thread = patch: LDG one tile ahead, centre, round / split, write the B operand, |z|^2,
// mask), warp 12 MMA issuer.  Two TMEM accumulator stages: the epilogue of tile i overlaps the MMAs of tile i+1.
#include "tc_head.cuh"
#include "tc_common.cuh"

namespace adlm {

namespace {

using namespace tc;

constexpr int TILE = 128;
constexpr float CENTRE = 0.5f;
constexpr int PUSH_THREADS = 416;

struct TcPushParams {
  const float* z;
  const uint32_t* mask;
  const float* P;
  const int32_t* proto_class;
  unsigned long long* packed_min;
  unsigned long long patch_base;
  int K, HW, N;
  int Npad;       // N rounded up to 16: rows of the prototype operand
  int M;          // n_img * HW
  int num_tiles;
};

template <int KP, bool X3>
struct PushSmem {
  static constexpr int PARTS = X3 ? 2 : 1;
  // prototype operand part: [KP/4][Npad][4] + slack so that an M=128 block starting at row 128 stays in bounds
  static __host__ __device__ size_t pa_part(int Npad, int NB) { return (size_t)(KP / 4) * Npad * 16 + (size_t)(NB * 128 - Npad) * 16; }
  static __host__ __device__ size_t pa(int part, int Npad, int NB) { return (size_t)part * pa_part(Npad, NB); }
  static __host__ __device__ size_t zb(int part, int Npad, int NB) { return PARTS * pa_part(Npad, NB) + (size_t)part * KP * TILE * 4; }
  static __host__ __device__ size_t misc(int Npad, int NB) { return zb(PARTS, Npad, NB); }
  // x2 [2][128] f32, mask [2][128] u32, 8 barriers, tmem base
  static __host__ __device__ size_t total(int Npad, int NB) { return misc(Npad, NB) + 2 * 128 * 4 * 2 + 64 + 16; }
};

template <int KP, bool X3, int NB>
__global__ void __launch_bounds__(PUSH_THREADS, 1) k_tc_push(const TcPushParams p) {
  using L = PushSmem<KP, X3>;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int Npad = p.Npad, K = p.K, N = p.N, HW = p.HW;
  float* PA_hi = reinterpret_cast<float*>(smem + L::pa(0, Npad, NB));
  float* PA_lo = reinterpret_cast<float*>(smem + L::pa(X3 ? 1 : 0, Npad, NB));
  float* ZB_hi = reinterpret_cast<float*>(smem + L::zb(0, Npad, NB));
  float* ZB_lo = reinterpret_cast<float*>(smem + L::zb(X3 ? 1 : 0, Npad, NB));
  uint8_t* misc = smem + L::misc(Npad, NB);
  float* x2s = reinterpret_cast<float*>(misc);                       // [2][128]
  uint32_t* msk = reinterpret_cast<uint32_t*>(misc + 1024);          // [2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(misc + 2048);
  uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(bars + 8);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (warp == 12) tmem_alloc(smem_u32(tmem_base_s), 512);
  if (tid == 0) {
    mbar_init(smem_u32(&bars[0]), 128);   // full: operand tile written
    mbar_init(smem_u32(&bars[1]), 1);     // stage_free: MMAs reading the operand tile are done
    mbar_init(smem_u32(&bars[2]), 1);     // acc_full[0]
    mbar_init(smem_u32(&bars[3]), 1);     // acc_full[1]
    mbar_init(smem_u32(&bars[4]), 256);   // acc_empty[0]
    mbar_init(smem_u32(&bars[5]), 256);   // acc_empty[1]
    mbar_init(smem_u32(&bars[6]), 128);   // meta_full[0]: |z|^2 and masks of the slot written
    mbar_init(smem_u32(&bars[7]), 128);   // meta_full[1]
    mbar_init_fence();
  }
  const uint32_t bar_full = smem_u32(&bars[0]), bar_sfree = smem_u32(&bars[1]);
  const uint32_t bar_afull0 = smem_u32(&bars[2]), bar_aempty0 = smem_u32(&bars[4]), bar_meta0 = smem_u32(&bars[6]);
  const int iters = p.num_tiles > (int)blockIdx.x ? (p.num_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  // producer warps start fetching their first tile before the prototypes are staged
  const bool is_producer = warp >= 8 && warp < 12;
  const int cw = tid - 256;                   // producer: patch column of the tile
  float zr[KP];
  uint32_t mk = 0u;
  auto prefetch = [&](int it) {
    const int m = ((int)blockIdx.x + it * (int)gridDim.x) * TILE + cw;
    if (it < iters && m < p.M) {
      const int img = m / HW;
      const float* src = p.z + ((size_t)img * K) * HW + (m - img * HW);
      if (K == KP) {
#pragma unroll
        for (int k = 0; k < KP; ++k) { zr[k] = __ldg(src); src += HW; }
      } else {
#pragma unroll
        for (int k = 0; k < KP; ++k) zr[k] = (k < K) ? __ldg(src + (size_t)k * HW) : CENTRE;
      }
      mk = __ldg(p.mask + m);
    } else {
#pragma unroll
      for (int k = 0; k < KP; ++k) zr[k] = CENTRE;
      mk = 0u;
    }
  };
  if (is_producer) prefetch(0);

  stage_rows_kmajor<KP, X3>(PA_hi, PA_lo, p.P, N, K, Npad, CENTRE, tid, PUSH_THREADS);
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem_base = *tmem_base_s;

  if (is_producer) {
    // ================= producer: thread = patch =================
    for (int it = 0; it < iters; ++it) {
      const int a = it & 1, j = it >> 1;
      if (it > 0) mbar_wait(bar_sfree, (uint32_t)(it - 1) & 1u);              // operand tile consumed
      if (j > 0) mbar_wait(bar_aempty0 + a * 8, (uint32_t)(j - 1) & 1u);      // x2 / mask slot drained
      float x2 = 0.f;
#pragma unroll
      for (int k4 = 0; k4 < KP / 4; ++k4) {
        float hi[4], lo[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float v = zr[k4 * 4 + q] - CENTRE;
          hi[q] = rna_tf32_mask(v);
          lo[q] = X3 ? rna_tf32_mask(v - hi[q]) : 0.f;
          const float u = X3 ? v : hi[q];
          x2 = fmaf(u, u, x2);
        }
        *reinterpret_cast<float4*>(ZB_hi + ((size_t)k4 * TILE + cw) * 4) = make_float4(hi[0], hi[1], hi[2], hi[3]);
        if (X3) *reinterpret_cast<float4*>(ZB_lo + ((size_t)k4 * TILE + cw) * 4) = make_float4(lo[0], lo[1], lo[2], lo[3]);
      }
      x2s[a * 128 + cw] = x2;
      msk[a * 128 + cw] = mk;
      fence_async_smem();
      mbar_arrive(bar_full);
      mbar_arrive(bar_meta0 + a * 8);
      prefetch(it + 1);
    }
  } else if (warp == 12) {
    // ================= MMA issuer =================
    if (lane == 0) {
      const uint32_t idesc = make_idesc_tf32(128, TILE);
      const int ksteps = (K + 7) >> 3;
      const uint32_t a_hi = smem_u32(PA_hi), a_lo = smem_u32(PA_lo), b_hi = smem_u32(ZB_hi), b_lo = smem_u32(ZB_lo);
      for (int it = 0; it < iters; ++it) {
        const int a = it & 1, j = it >> 1;
        mbar_wait(bar_full, (uint32_t)it & 1u);
        if (j > 0) mbar_wait(bar_aempty0 + a * 8, (uint32_t)(j - 1) & 1u);
        fence_after();
#pragma unroll
        for (int blk = 0; blk < NB; ++blk) {
          const uint32_t d = tmem_base + a * 256 + blk * 128;
          uint32_t acc = 0u;
          for (int ks = 0; ks < ksteps; ++ks) {
            const uint32_t ao = blk * 128 * 16 + ks * 2 * Npad * 16, bo = ks * 2 * TILE * 16;
            if (X3) {
              mma_ss(d, make_desc(a_hi + ao, Npad * 16, 128), make_desc(b_lo + bo, TILE * 16, 128), idesc, acc);
              mma_ss(d, make_desc(a_lo + ao, Npad * 16, 128), make_desc(b_hi + bo, TILE * 16, 128), idesc, 1u);
              acc = 1u;
            }
            mma_ss(d, make_desc(a_hi + ao, Npad * 16, 128), make_desc(b_hi + bo, TILE * 16, 128), idesc, acc);
            acc = 1u;
          }
        }
        mma_commit(bar_sfree);
        mma_commit(bar_afull0 + a * 8);
      }
    }
  } else {
    // ================= epilogue: thread = prototype =================
    // NB == 2: warps 0-3 own prototypes 0-127, warps 4-7 prototypes 128-255, all 128 columns each.
    // NB == 1: all eight warps own prototypes 0-127; warps 4-7 take columns 64-127.
    const int q = warp & 3, grp = warp >> 2;
    const int blk = NB == 2 ? grp : 0;
    const int c_begin = NB == 2 ? 0 : grp * 64, c_end = NB == 2 ? TILE : c_begin + 64;
    const int n = blk * 128 + q * 32 + lane;
    const uint32_t lane_field = (uint32_t)(q * 32) << 16;
    float p2 = 0.f;
    uint32_t cbit = 0u;
    if (n < N) {
      for (int k4 = 0; k4 < KP / 4; ++k4) {
        float4 v = *reinterpret_cast<const float4*>(PA_hi + ((size_t)k4 * Npad + n) * 4);
        if (X3) {
          const float4 l4 = *reinterpret_cast<const float4*>(PA_lo + ((size_t)k4 * Npad + n) * 4);
          v.x += l4.x; v.y += l4.y; v.z += l4.z; v.w += l4.w;
        }
        p2 = fmaf(v.x, v.x, p2); p2 = fmaf(v.y, v.y, p2); p2 = fmaf(v.z, v.z, p2); p2 = fmaf(v.w, v.w, p2);
      }
      const int pc = p.proto_class[n];
      if (pc >= 0 && pc < 32) cbit = 1u << pc;
    }
    const bool warp_has_work = (blk * 128 + q * 32) < N;     // warps whose 32 prototypes are all padding skip the math
    float best = __int_as_float(0x7f800000);                 // +inf
    int best_m = -1;
    for (int it = 0; it < iters; ++it) {
      const int a = it & 1, j = it >> 1;
      const int m0 = ((int)blockIdx.x + it * (int)gridDim.x) * TILE;
      mbar_wait(bar_meta0 + a * 8, (uint32_t)j & 1u);
      mbar_wait(bar_afull0 + a * 8, (uint32_t)j & 1u);
      fence_after();
      if (warp_has_work) {
        const uint32_t t0 = tmem_base + a * 256 + blk * 128 + lane_field;
        const float* x2a = x2s + a * 128;
        const uint32_t* mka = msk + a * 128;
        uint32_t ra[16], rb[16];
        auto chunk = [&](int c0, const uint32_t (&r)[16]) {
#pragma unroll
          for (int v4 = 0; v4 < 4; ++v4) {
            const float4 xq = *reinterpret_cast<const float4*>(x2a + c0 + v4 * 4);
            const uint4 mq = *reinterpret_cast<const uint4*>(mka + c0 + v4 * 4);
            const float xs[4] = {xq.x, xq.y, xq.z, xq.w};
            const uint32_t ms[4] = {mq.x, mq.y, mq.z, mq.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int i = v4 * 4 + e;
              const float d = fmaxf(fmaf(-2.f, __uint_as_float(r[i]), xs[e] + p2), 0.f);
              if ((ms[e] & cbit) != 0u && d < best) { best = d; best_m = m0 + c0 + i; }
            }
          }
        };
        tmem_ld16(t0 + c_begin, ra);
        for (int c0 = c_begin; c0 < c_end; c0 += 32) {
          tmem_ld_wait();
          tmem_ld16(t0 + c0 + 16, rb);
          chunk(c0, ra);
          tmem_ld_wait();
          if (c0 + 32 < c_end) tmem_ld16(t0 + c0 + 32, ra);
          chunk(c0 + 16, rb);
        }
      }
      fence_before();
      mbar_arrive(bar_aempty0 + a * 8);
    }
    // padding columns of the last tile (m >= M) carry mask 0 and never win
    if (n < N && best_m >= 0) {
      const unsigned long long key =
          ((unsigned long long)__float_as_uint(best) << 32) | (p.patch_base + (unsigned long long)best_m);
      atomicMin(p.packed_min + n, key);
    }
  }
  fence_before();
  __syncthreads();
  if (warp == 12) tmem_dealloc(tmem_base, 512);
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
  const size_t bytes = L::total(prm.Npad, NB);
  if (bytes > 227 * 1024) {
    set_error("tc_push_scan: %zu bytes of shared memory needed (> 227 KB) for K=%d N=%d", bytes, prm.K, prm.N);
    return ADLM_E_UNSUPPORTED;
  }
  ADLM_CUDA(cudaFuncSetAttribute(k_tc_push<KP, X3, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  const int grid = prm.num_tiles < sm_count_push() ? prm.num_tiles : sm_count_push();
  KernelTimer timer(ADLM_K_PUSH_SCAN, st);
  k_tc_push<KP, X3, NB><<<grid, PUSH_THREADS, bytes, st>>>(prm);
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

}  // namespace

int tc_push_scan(const float* z, int n_img, int K, int HW, const uint32_t* mask, const float* P, int N,
                 const int32_t* proto_class, long long patch_base, int prec_mode,
                 unsigned long long* packed_min, cudaStream_t st) {
  const long long M = (long long)n_img * HW;
  // shapes beyond the tensor-core kernel's operand budget take the fp32 CUDA-core scan: same search, same keys
  if (K > 64 || N > 256 || M >= (1ll << 31) - 256)
    return simt_push_scan(z, n_img, K, HW, mask, P, N, proto_class, patch_base, packed_min, st);
  TcPushParams prm;
  prm.z = z; prm.mask = mask; prm.P = P; prm.proto_class = proto_class; prm.packed_min = packed_min;
  prm.patch_base = (unsigned long long)patch_base;
  prm.K = K; prm.HW = HW; prm.N = N;
  prm.Npad = (N + 15) & ~15;
  prm.M = (int)M;
  prm.num_tiles = (int)((M + TILE - 1) / TILE);
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
