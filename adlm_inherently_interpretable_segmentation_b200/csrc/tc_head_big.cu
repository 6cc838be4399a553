// Prototype-head forward for LARGE heads on the tensor cores (ADLM_PREC_TF32): N <= 1024 prototypes, K <= 256 depth,
// C <= 112 classes — BASELINE configs[4] (1000 prototypes x 256-d, 100 classes on 1024 x 2048 maps), the shapes the
// warp-specialised kernel of tc_head_ws.cu cannot hold in shared memory.  Same mathematics (model.py:203-283).
//
// A 128-pixel tile keeps its features resident in shared memory as the A operand (K-major, all of K: 128 KB at
// K = 256) and sweeps the prototypes in blocks of 128:
//   MMA1  D1[b] = A . Pblk^T        32 x (M128 N128 K8); the block's prototypes stream through a 4-slot ring of 16 KB
//                                   chunks [32 depth x 128 prototypes] fetched by cp.async.bulk (TMA bulk copy) from a
//                                   pre-staged copy of P that already has the operand layout (centred, rounded)
//   E1    d = max(|z|^2 - 2 acc + |p|^2, 0) -> s = log2(1 + (1-eps)/(d+eps)) -> back into D1[b] (TMEM A operand)
//   MMA2  D2[t&1] += S . Wblk^T     16 x (M128 N=Cpad K8), W chunks [32 prototypes x Cpad] through a 2-slot ring,
//                                   accumulated over the prototype blocks: the last layer never sees an M x N map
//   E2    logits / per-pixel argmax from D2 after the last block
// D1 is double-buffered (MMA1 of block j+1 runs under the epilogue of block j), D2 by tile parity.
// Warps: 0-7 epilogue, 8-11 producers (thread = pixel: coalesced LDG of the NCHW tile, centre, round, |z|^2),
// 12 MMA issue (whole warp around elect.sync), 13 prototype-chunk loader, 14 last-layer-chunk loader.
#include "tc_head.cuh"
#include "tc_common.cuh"

namespace adlm {

namespace {

using namespace tc;

constexpr int TILE = 128, NBLK = 128, KCH = 32, BIG_THREADS = 480;
constexpr int P_SLOTS = 4, W_SLOTS = 2;
constexpr float LN2 = 0.6931471805599453f;
constexpr float CENTRE = 0.5f;
constexpr int P_CHUNK_FLOATS = (KCH / 4) * NBLK * 4;            // 4096 floats = 16 KB

struct BigParams {
  const float* z;
  const float* Pst;       // staged prototypes [nblk][KC][8][128][4]
  const float* p2;        // [Npad] |p|^2 of the rounded, centred prototypes
  const float* Wst;       // staged last layer [nblk][4][8][Cpad][4], ln 2 folded in
  float* logits;          // [M, C]
  float* dist;            // [B, N, HW] | null
  int32_t* argmax;        // [M] | null
  int B, K, HW, N, C;
  int KC;                 // depth chunks of 32
  int nblk;               // prototype blocks of 128
  int Cpad;               // classes rounded up to 16
  float eps;
  int act_mode;
  long long M;
  int num_tiles;
};

// ---- one-time staging of P and W into the operand layouts (a few microseconds: ~1.5 MB) ----
__global__ void __launch_bounds__(256) k_big_stage(const float* __restrict__ P, const float* __restrict__ W, int N, int K,
                                                   int C, int KC, int nblk, int Cpad, float wscale,
                                                   float* __restrict__ Pst, float* __restrict__ p2,
                                                   float* __restrict__ Wst) {
  const int Npad = nblk * NBLK;
  const long long nP = (long long)nblk * KC * 8 * NBLK;          // float4 groups of Pst
  const long long nW = (long long)nblk * 4 * 8 * Cpad;           // float4 groups of Wst
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long f = tid; f < nP; f += stride) {
    const int row = (int)(f % NBLK);
    const int k4 = (int)((f / NBLK) % 8);
    const int kc = (int)((f / (NBLK * 8)) % KC);
    const int blk = (int)(f / ((long long)NBLK * 8 * KC));
    const int n = blk * NBLK + row;
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int k = kc * KCH + k4 * 4 + e;
      v[e] = (n < N && k < K) ? rna_tf32_mask(__ldg(P + (size_t)n * K + k) - CENTRE) : 0.f;
    }
    reinterpret_cast<float4*>(Pst)[f] = make_float4(v[0], v[1], v[2], v[3]);
  }
  for (long long f = tid; f < nW; f += stride) {
    const int c = (int)(f % Cpad);
    const int k4 = (int)((f / Cpad) % 8);
    const int kc2 = (int)((f / (Cpad * 8)) % 4);
    const int blk = (int)(f / ((long long)Cpad * 8 * 4));
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int n = blk * NBLK + kc2 * 32 + k4 * 4 + e;
      v[e] = (c < C && n < N) ? rna_tf32_mask(wscale * __ldg(W + (size_t)c * N + n)) : 0.f;
    }
    reinterpret_cast<float4*>(Wst)[f] = make_float4(v[0], v[1], v[2], v[3]);
  }
  for (long long n = tid; n < Npad; n += stride) {
    float a = 0.f;
    if (n < N)
      for (int k = 0; k < K; ++k) {          // same rounding and the same k order as the operand
        const float x = rna_tf32_mask(__ldg(P + (size_t)n * K + k) - CENTRE);
        a = fmaf(x, x, a);
      }
    p2[n] = a;
  }
}

// barrier indices
constexpr int B_AFULL = 0;               // A tile + |z|^2 written (128 producer threads), once per tile
constexpr int B_AFREE = 1;               // last MMA1 of the tile has read A (commit)
constexpr int B_D1FULL = 2;              // [2] D1[b] ready (commit)
constexpr int B_SREADY = 4;              // [2] similarities back in D1[b] (256 epilogue threads)
constexpr int B_D2FULL = 6;              // [2] D2[tile parity] complete (commit)
constexpr int B_PFULL = 8;               // [P_SLOTS] prototype chunk landed (expect_tx)
constexpr int B_PEMPTY = 12;             // [P_SLOTS] MMAs of the chunk done (commit)
constexpr int B_WFULL = 16;              // [W_SLOTS]
constexpr int B_WEMPTY = 18;             // [W_SLOTS]
constexpr int NBARS = 20;

__device__ __forceinline__ void bulk_load(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_smem), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

struct BigSmem {
  // A [KPAD/4][128][4] | P ring | W ring | p2 [Npad] | x2 [2][128] | barriers | tmem base
  static __host__ __device__ size_t a() { return 0; }
  static __host__ __device__ size_t pring(int KC) { return (size_t)KC * KCH * TILE * 4; }
  static __host__ __device__ size_t wring(int KC) { return pring(KC) + (size_t)P_SLOTS * P_CHUNK_FLOATS * 4; }
  static __host__ __device__ size_t p2(int KC, int Cpad) { return wring(KC) + (size_t)W_SLOTS * 32 * Cpad * 4; }
  static __host__ __device__ size_t x2(int KC, int Cpad, int Npad) { return p2(KC, Cpad) + (size_t)Npad * 4; }
  static __host__ __device__ size_t bars(int KC, int Cpad, int Npad) { return x2(KC, Cpad, Npad) + 2 * TILE * 4; }
  static __host__ __device__ size_t total(int KC, int Cpad, int Npad) { return bars(KC, Cpad, Npad) + NBARS * 8 + 16; }
};

__global__ void __launch_bounds__(BIG_THREADS, 1) k_tc_fwd_big(const BigParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int KC = p.KC, nblk = p.nblk, Cpad = p.Cpad, Npad = nblk * NBLK, K = p.K, N = p.N, C = p.C, HW = p.HW;
  float* A = reinterpret_cast<float*>(smem + BigSmem::a());
  float* Pr = reinterpret_cast<float*>(smem + BigSmem::pring(KC));
  float* Wr = reinterpret_cast<float*>(smem + BigSmem::wring(KC));
  float* p2s = reinterpret_cast<float*>(smem + BigSmem::p2(KC, Cpad));
  float* x2s = reinterpret_cast<float*>(smem + BigSmem::x2(KC, Cpad, Npad));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + BigSmem::bars(KC, Cpad, Npad));
  uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(bars + NBARS);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int Mi = (int)p.M;
  const int iters = p.num_tiles > (int)blockIdx.x ? (p.num_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  auto tile_of = [&](int it) { return (int)blockIdx.x + it * (int)gridDim.x; };
  const uint32_t w_chunk_bytes = (uint32_t)(32 * Cpad * 4);

  if (warp == 12) tmem_alloc(smem_u32(tmem_base_s), 512);
  if (tid == 0) {
    mbar_init(smem_u32(&bars[B_AFULL]), 128);
    mbar_init(smem_u32(&bars[B_AFREE]), 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&bars[B_D1FULL + i]), 1);
      mbar_init(smem_u32(&bars[B_SREADY + i]), 256);
      mbar_init(smem_u32(&bars[B_D2FULL + i]), 1);
    }
    for (int i = 0; i < P_SLOTS; ++i) {
      mbar_init(smem_u32(&bars[B_PFULL + i]), 1);
      mbar_init(smem_u32(&bars[B_PEMPTY + i]), 1);
    }
    for (int i = 0; i < W_SLOTS; ++i) {
      mbar_init(smem_u32(&bars[B_WFULL + i]), 1);
      mbar_init(smem_u32(&bars[B_WEMPTY + i]), 1);
    }
    mbar_init_fence();
  }
  for (int n = tid; n < Npad; n += BIG_THREADS) p2s[n] = __ldg(p.p2 + n);
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem_base = *tmem_base_s;
  const uint32_t bar0 = smem_u32(&bars[0]);
  auto bar = [&](int idx) { return bar0 + (uint32_t)idx * 8u; };
  const uint32_t tD1 = tmem_base, tD2 = tmem_base + 2 * NBLK;     // D1[b] at b*128, D2[t&1] at 256 + (t&1)*Cpad

  if (warp >= 8 && warp < 12) {
    // ================= producers: thread = pixel =================
    const int cw = tid - 256;
    for (int it = 0; it < iters; ++it) {
      const int m = tile_of(it) * TILE + cw;
      const bool valid = m < Mi;
      const int b = valid ? m / HW : 0;
      const float* src = p.z + ((size_t)b * K) * HW + (valid ? m - b * HW : 0);
      float zr[KCH];
      auto load = [&](int kc) {
#pragma unroll
        for (int k = 0; k < KCH; ++k) {
          const int kk = kc * KCH + k;
          zr[k] = (valid && kk < K) ? __ldg(src + (size_t)kk * HW) : CENTRE;
        }
      };
      load(0);                                                   // in flight while the previous tile still owns A
      if (it > 0) mbar_wait(bar(B_AFREE), (uint32_t)(it - 1) & 1u);
      float x2 = 0.f;
      for (int kc = 0; kc < KC; ++kc) {
        float v[KCH];
#pragma unroll
        for (int k = 0; k < KCH; ++k) {
          v[k] = rna_tf32_mask(zr[k] - CENTRE);
          x2 = fmaf(v[k], v[k], x2);
        }
        if (kc + 1 < KC) load(kc + 1);
#pragma unroll
        for (int k4 = 0; k4 < KCH / 4; ++k4)
          *reinterpret_cast<float4*>(A + ((size_t)(kc * (KCH / 4) + k4) * TILE + cw) * 4) =
              make_float4(v[k4 * 4], v[k4 * 4 + 1], v[k4 * 4 + 2], v[k4 * 4 + 3]);
      }
      x2s[(it & 1) * TILE + cw] = x2;
      fence_async_smem();
      mbar_arrive(bar(B_AFULL));
    }
  } else if (warp == 13) {
    // ================= prototype-chunk loader (one thread, TMA bulk copies) =================
    if (lane == 0) {
      uint32_t i = 0;
      for (int it = 0; it < iters; ++it)
        for (int blk = 0; blk < nblk; ++blk)
          for (int kc = 0; kc < KC; ++kc, ++i) {
            const uint32_t slot = i % P_SLOTS, use = i / P_SLOTS;
            if (use > 0) mbar_wait(bar(B_PEMPTY + slot), (use - 1) & 1u);
            bulk_load(smem_u32(Pr + (size_t)slot * P_CHUNK_FLOATS), p.Pst + ((size_t)blk * KC + kc) * P_CHUNK_FLOATS,
                      P_CHUNK_FLOATS * 4, bar(B_PFULL + slot));
          }
    }
  } else if (warp == 14) {
    // ================= last-layer-chunk loader =================
    if (lane == 0) {
      uint32_t i = 0;
      for (int it = 0; it < iters; ++it)
        for (int blk = 0; blk < nblk; ++blk)
          for (int kc2 = 0; kc2 < 4; ++kc2, ++i) {
            const uint32_t slot = i % W_SLOTS, use = i / W_SLOTS;
            if (use > 0) mbar_wait(bar(B_WEMPTY + slot), (use - 1) & 1u);
            bulk_load(smem_u32(Wr + (size_t)slot * 32 * Cpad), p.Wst + ((size_t)blk * 4 + kc2) * 32 * Cpad, w_chunk_bytes,
                      bar(B_WFULL + slot));
          }
    }
  } else if (warp == 12) {
    // ================= MMA issue (whole warp, one elected lane per instruction) =================
    const uint32_t idesc1 = make_idesc_tf32(TILE, NBLK), idesc2 = make_idesc_tf32(TILE, Cpad);
    const uint32_t a_base = smem_u32(A), p_base = smem_u32(Pr), w_base = smem_u32(Wr);
    uint32_t pi = 0, wi = 0;                                    // chunks consumed so far
    uint32_t d1cnt[2] = {0u, 0u};
    auto mma1 = [&](int blk, bool last_of_tile) {
      const int b = blk & 1;
      for (int kc = 0; kc < KC; ++kc, ++pi) {
        const uint32_t slot = pi % P_SLOTS, use = pi / P_SLOTS;
        mbar_wait(bar(B_PFULL + slot), use & 1u);
        fence_after();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int ks = kc * 4 + j;
          mma_ss_w(tD1 + b * NBLK, make_desc(a_base + ks * 2 * TILE * 16, TILE * 16, 128),
                   make_desc(p_base + slot * (P_CHUNK_FLOATS * 4) + j * 2 * NBLK * 16, NBLK * 16, 128), idesc1,
                   ks > 0 ? 1u : 0u);
        }
        mma_commit_w(bar(B_PEMPTY + slot));
      }
      mma_commit_w(bar(B_D1FULL + b));
      if (last_of_tile) mma_commit_w(bar(B_AFREE));
    };
    auto mma2 = [&](int it, int blk) {
      const int b = blk & 1, tp = it & 1;
      mbar_wait(bar(B_SREADY + b), d1cnt[b] & 1u);
      ++d1cnt[b];
      fence_after();
      for (int kc2 = 0; kc2 < 4; ++kc2, ++wi) {
        const uint32_t slot = wi % W_SLOTS, use = wi / W_SLOTS;
        mbar_wait(bar(B_WFULL + slot), use & 1u);
        fence_after();
#pragma unroll
        for (int j = 0; j < 4; ++j)
          mma_ts_w(tD2 + tp * Cpad, tD1 + b * NBLK + (kc2 * 4 + j) * 8,
                   make_desc(w_base + slot * w_chunk_bytes + j * 2 * Cpad * 16, Cpad * 16, 128), idesc2,
                   (blk > 0 || kc2 > 0 || j > 0) ? 1u : 0u);
        mma_commit_w(bar(B_WEMPTY + slot));
      }
      if (blk == nblk - 1) mma_commit_w(bar(B_D2FULL + tp));
    };
    for (int it = 0; it < iters; ++it) {
      mbar_wait(bar(B_AFULL), (uint32_t)it & 1u);
      fence_after();
      mma1(0, nblk == 1);
      for (int blk = 0; blk < nblk; ++blk) {
        if (blk + 1 < nblk) mma1(blk + 1, blk + 1 == nblk - 1);
        mma2(it, blk);
      }
    }
  } else {
    // ================= epilogue warps =================
    const int q = warp & 3, half = warp >> 2;
    const int row = q * 32 + lane;
    const uint32_t lane_field = (uint32_t)(q * 32) << 16;
    const bool act_log = p.act_mode == ADLM_ACT_LOG;
    const float eps = p.eps, one_m_eps = 1.f - p.eps;
    uint32_t d1cnt[2] = {0u, 0u};

    auto epilogue2 = [&](int t) {
      const int tp = t & 1;
      const int m = tile_of(t) * TILE + row;
      mbar_wait(bar(B_D2FULL + tp), (uint32_t)(t >> 1) & 1u);
      fence_after();
      const uint32_t d2 = tD2 + tp * Cpad + lane_field;
      float bv = -INFINITY;
      int best = 0;
      float* dst = p.logits + (size_t)m * C;
      const bool vec = (C & 3) == 0 && (reinterpret_cast<uintptr_t>(p.logits) & 15u) == 0;
      for (int c0 = 0; c0 < Cpad; c0 += 16) {
        uint32_t r[16];
        tmem_ld16(d2 + c0, r);
        tmem_ld_wait();
        if (m < Mi) {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float v = __uint_as_float(r[i]);
            if (c0 + i < C && v > bv) { bv = v; best = c0 + i; }
          }
          if (vec) {
#pragma unroll
            for (int i4 = 0; i4 < 4; ++i4)
              if (c0 + i4 * 4 < C)
                *reinterpret_cast<float4*>(dst + c0 + i4 * 4) =
                    make_float4(__uint_as_float(r[i4 * 4]), __uint_as_float(r[i4 * 4 + 1]), __uint_as_float(r[i4 * 4 + 2]),
                                __uint_as_float(r[i4 * 4 + 3]));
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (c0 + i < C) dst[c0 + i] = __uint_as_float(r[i]);
          }
        }
      }
      if (m < Mi && p.argmax != nullptr) p.argmax[m] = best;
    };

    for (int it = 0; it < iters; ++it) {
      const int m = tile_of(it) * TILE + row;
      const bool valid = m < Mi;
      const int b = valid ? m / HW : 0;
      const int pix = valid ? m - b * HW : 0;
      // |z|^2 of the tile is read after the tile's first D1 is complete (the MMA thread issued it after it observed
      // the producers' barrier).  The epilogue does not wait on that barrier itself: with one or two prototype blocks
      // it could complete twice before a late thread looks, and a parity wait cannot tell that apart.
      float x2 = 0.f;
      for (int blk = 0; blk < nblk; ++blk) {
        const int db = blk & 1;
        mbar_wait(bar(B_D1FULL + db), d1cnt[db] & 1u);
        ++d1cnt[db];
        fence_after();
        if (blk == 0) x2 = x2s[(it & 1) * TILE + row];
        const uint32_t t1 = tD1 + db * NBLK + lane_field;
#pragma unroll 1
        for (int ch = 0; ch < 4; ++ch) {
          const int c0 = half * 64 + ch * 16, n0 = blk * NBLK + c0;
          uint32_t r[16], sb[16];
          tmem_ld16(t1 + c0, r);
          float xp[16];
#pragma unroll
          for (int v4 = 0; v4 < 4; ++v4) {
            const float4 pq = *reinterpret_cast<const float4*>(p2s + n0 + v4 * 4);
            xp[v4 * 4] = x2 + pq.x; xp[v4 * 4 + 1] = x2 + pq.y; xp[v4 * 4 + 2] = x2 + pq.z; xp[v4 * 4 + 3] = x2 + pq.w;
          }
          tmem_ld_wait();
          float dv[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float d = fmaxf(fmaf(-2.f, __uint_as_float(r[i]), xp[i]), 0.f);
            dv[i] = d;
            const float s = act_log ? lg2_approx(fmaf(one_m_eps, rcp_approx(d + eps), 1.f)) : -d;
            sb[i] = __float_as_uint(rna_tf32_fast(s));
          }
          tmem_st16(t1 + c0, sb);
          if (p.dist != nullptr && valid) {
            float* dd = p.dist + ((size_t)b * N + n0) * HW + pix;
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (n0 + i < N) dd[(size_t)i * HW] = dv[i];
          }
        }
        tmem_st_wait();
        fence_before();
        mbar_arrive(bar(B_SREADY + db));
        // logits of the previous tile (its D2 is complete since its last block): the two halves alternate
        if (blk == 0 && it > 0 && ((it - 1) & 1) == half) epilogue2(it - 1);
      }
    }
    if (iters > 0 && ((iters - 1) & 1) == half) epilogue2(iters - 1);
  }
  fence_before();
  __syncthreads();
  if (warp == 12) tmem_dealloc(tmem_base, 512);
}

int g_sms_big = 0;
int sm_count_big() {
  if (g_sms_big == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sms_big, cudaDevAttrMultiProcessorCount, dev);
    if (g_sms_big <= 0) g_sms_big = 148;
  }
  return g_sms_big;
}

}  // namespace

bool tc_head_big_supported(const HeadShape& s) { return s.K <= 256 && s.N <= 1024 && s.C <= 112; }

size_t tc_head_big_workspace_bytes(int K, int N, int C) {
  const int KC = (K + KCH - 1) / KCH, nblk = (N + NBLK - 1) / NBLK, Cpad = (C + 15) & ~15;
  return align_up((size_t)nblk * KC * P_CHUNK_FLOATS * 4, 256) + align_up((size_t)nblk * NBLK * 4, 256) +
         align_up((size_t)nblk * 4 * 32 * Cpad * 4, 256);
}

int tc_head_fwd_big(const HeadShape& s, const float* z, const float* P, const float* W, float* logits, float* dist,
                    int32_t* argmax, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  if (!tc_head_big_supported(s)) {
    set_error("adlm_head_fwd_big: needs K <= 256, N <= 1024, C <= 112 (got K=%d N=%d C=%d); use ADLM_PREC_FP32", s.K, s.N, s.C);
    return ADLM_E_UNSUPPORTED;
  }
  const size_t need = tc_head_big_workspace_bytes(s.K, s.N, s.C);
  if (workspace == nullptr || workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 255u) != 0) {
    set_error("adlm_head_fwd_big: workspace of %zu bytes, 256-byte aligned, needed (got %zu)", need, workspace_bytes);
    return ADLM_E_WORKSPACE;
  }
  BigParams prm;
  prm.KC = (s.K + KCH - 1) / KCH;
  prm.nblk = (s.N + NBLK - 1) / NBLK;
  prm.Cpad = (s.C + 15) & ~15;
  char* w8 = static_cast<char*>(workspace);
  float* Pst = reinterpret_cast<float*>(w8);
  float* p2 = reinterpret_cast<float*>(w8 + align_up((size_t)prm.nblk * prm.KC * P_CHUNK_FLOATS * 4, 256));
  float* Wst = reinterpret_cast<float*>(reinterpret_cast<char*>(p2) + align_up((size_t)prm.nblk * NBLK * 4, 256));
  prm.z = z; prm.Pst = Pst; prm.p2 = p2; prm.Wst = Wst;
  prm.logits = logits; prm.dist = dist; prm.argmax = argmax;
  prm.B = s.B; prm.K = s.K; prm.HW = s.HW; prm.N = s.N; prm.C = s.C;
  prm.eps = s.eps; prm.act_mode = s.act_mode;
  prm.M = (long long)s.B * s.HW;
  prm.num_tiles = (int)((prm.M + TILE - 1) / TILE);
  const size_t bytes = BigSmem::total(prm.KC, prm.Cpad, prm.nblk * NBLK);
  if (bytes > 227 * 1024) {
    set_error("adlm_head_fwd_big: %zu bytes of shared memory needed (> 227 KB)", bytes);
    return ADLM_E_UNSUPPORTED;
  }
  k_big_stage<<<256, 256, 0, st>>>(P, W, s.N, s.K, s.C, prm.KC, prm.nblk, prm.Cpad,
                                   s.act_mode == ADLM_ACT_LOG ? LN2 : 1.f, Pst, p2, Wst);
  ADLM_LAUNCH_CHECK();
  ADLM_CUDA(cudaFuncSetAttribute(k_tc_fwd_big, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  const int grid = prm.num_tiles < sm_count_big() ? prm.num_tiles : sm_count_big();
  KernelTimer timer(ADLM_K_HEAD_FWD, st);
  k_tc_fwd_big<<<grid, BIG_THREADS, bytes, st>>>(prm);
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

}  // namespace adlm
