// Prototype-head forward for LARGE heads on the tensor cores (ADLM_PREC_TF32): N <= 1024 prototypes, K <= 256 depth,
// C <= 112 classes — BASELINE configs[4] (1000 prototypes x 256-d, 100 classes on 1024 x 2048 maps), the shapes the
// warp-specialised kernel of tc_head_ws.cu cannot hold in shared memory.  Same mathematics (model.py:203-283).
//
// The kernel runs on CTA PAIRS (cluster of two CTAs on one TPC, tcgen05 cta_group::2).  Each CTA of a pair owns one
// 128-pixel tile whose features stay resident in its shared memory as the A operand (K-major, all of K: 128 KB at
// K = 256); the pair sweeps the prototypes in blocks of 128 with M = 256 MMAs issued by the leader CTA:
//   MMA1  D1[b] = A . Pblk^T        32 x (M256 N128 K8); each CTA streams only HALF of the block's prototypes (64 rows)
//                                   through an 8-slot ring of 8 KB chunks [32 depth x 64 prototypes] fetched by
//                                   cp.async.bulk (TMA bulk copy) from a pre-staged copy of P in operand layout
//   E1    d = max(|z|^2 - 2 acc + |p|^2, 0) -> s = log2(1 + (1-eps)/(d+eps)) -> back into D1[b] (TMEM A operand)
//   MMA2  D2[t&1] += S . Wblk^T     16 x (M256 N=Cpad K8), each CTA streams Cpad/2 class rows per W chunk through a
//                                   2-slot ring; accumulated over the prototype blocks: no M x N map in memory
//   E2    logits / per-pixel argmax from D2 after the last block
// Why pairs: with one CTA per tile every 128 pixels pulled all of P and W (1.46 MB at cfg5) from L2, 6 GB per
// 512 x 1024 map; with the MMAs and the epilogue switched off the kernel still took 0.79 of its 1.15 ms, i.e. it sat
// on the ~7.6 TB/s L2 -> SM ceiling at 42 % of the TF32 matmul peak (profiles/r2_cfg5_experiments.log).
// A pair shares each B operand, which halves that traffic.
// D1 is double-buffered (MMA1 of block j+1 runs under the epilogue of block j), D2 by tile parity.
// Warps (per CTA): 0-7 epilogue, 8-11 producers (thread = pixel: coalesced LDG of the NCHW tile, centre, round, |z|^2),
// 12 MMA issue in the leader (whole warp around elect.sync) / in the peer a forwarder that tells the leader when the
// peer's chunks have landed, 13 prototype-chunk loader, 14 last-layer-chunk loader.
// Barriers: "full" events of the peer reach the LEADER's barriers (producer and epilogue warps arrive remotely, the
// forwarder relays the TMA completions); "free" events are tcgen05.commit multicasts onto both CTAs' barriers.
#include <cstdio>
#include <cstdlib>
#include "tc_head.cuh"
#include "tc_common.cuh"

namespace adlm {

namespace {

using namespace tc;

constexpr int TILE = 128, NBLK = 128, KCH = 32, BIG_THREADS = 480;
constexpr int P_SLOTS_MAX = 16, W_SLOTS = 2;
constexpr int NHALF = NBLK / 2;                                 // prototype rows of a block held by one CTA
constexpr float LN2 = 0.6931471805599453f;
constexpr float CENTRE = 0.5f;
constexpr int P_CHUNK_FLOATS = (KCH / 4) * NHALF * 4;           // 2048 floats = 8 KB per CTA
// Chunks travel and are released in GROUPS: one bulk copy, one barrier round and one commit per PG prototype chunks
// (WG last-layer chunks).  The fixed cost of a ring step (two barrier waits, fence, commit: ~230 cycles that do not
// overlap the MMAs) is then paid per 8 MMAs instead of per 4; the ring's depth itself does not matter (2 slots run as
// fast as 16, profiles/r2_cfg5_experiments.log).
// PG = 4 (16 MMAs per ring step, two slots of 32 KB) when the depth is a multiple of 128, else 2: 2.93 -> 2.75 ms at
// cfg5.  The depth is padded to a multiple of PG chunks, so PG = 4 would double the work of a 64-deep head.
constexpr int WG = 2;
__host__ __device__ constexpr int pick_pg(int K) { return ((K + KCH - 1) / KCH) % 4 == 0 ? 4 : 2; }

struct BigParams {
  const float* z;
  const float* Pst;       // staged prototypes [nblk][KC/PG][2 halves][PG][8][64][4]
  const float* p2;        // [Npad] |p|^2 of the rounded, centred prototypes
  const float* Wst;       // staged last layer [nblk][4/WG][2 halves][WG][8][Cpad/2][4], ln 2 folded in
  float* logits;          // [M, C]
  float* dist;            // [B, N, HW] | null
  int32_t* argmax;        // [M] | null
  int B, K, HW, N, C;
  int KC;                 // depth chunks of 32, rounded up to a multiple of PG (the padding is zeros on both sides)
  int PG;                 // prototype chunks per ring step (2 | 4)
  int nblk;               // prototype blocks of 128
  int Cpad;               // classes rounded up to 16
  float eps;
  int act_mode;
  long long M;
  int num_tiles;
  int p_slots;            // depth of the prototype-chunk ring (what shared memory leaves, at most P_SLOTS_MAX)
  int exp_flags;          // timing experiments (ADLM_BIG_EXP; results are wrong when set): 1 no MMA1, 2 no MMA2,
                          // 4 no epilogue arithmetic
};

// ---- one-time staging of P and W into the operand layouts (a few microseconds: ~1.5 MB) ----
__global__ void __launch_bounds__(256) k_big_stage(const float* __restrict__ P, const float* __restrict__ W, int N, int K,
                                                   int C, int KC, int PG, int nblk, int Cpad, float wscale,
                                                   float* __restrict__ Pst, float* __restrict__ p2,
                                                   float* __restrict__ Wst) {
  const int Npad = nblk * NBLK;
  const long long nP = (long long)nblk * KC * 8 * NBLK;          // float4 groups of Pst
  const long long nW = (long long)nblk * 4 * 8 * Cpad;           // float4 groups of Wst
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long f = tid; f < nP; f += stride) {
    const int row = (int)(f % NHALF);
    const int k4 = (int)((f / NHALF) % 8);
    const int g = (int)((f / (NHALF * 8)) % PG);
    const int half = (int)((f / (NHALF * 8 * PG)) % 2);
    const int kc = (int)((f / (NBLK * 8 * PG)) % (KC / PG)) * PG + g;
    const int blk = (int)(f / ((long long)NBLK * 8 * KC));
    const int n = blk * NBLK + half * NHALF + row;
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int k = kc * KCH + k4 * 4 + e;
      v[e] = (n < N && k < K) ? rna_tf32_mask(__ldg(P + (size_t)n * K + k) - CENTRE) : 0.f;
    }
    reinterpret_cast<float4*>(Pst)[f] = make_float4(v[0], v[1], v[2], v[3]);
  }
  const int Ch = Cpad / 2;
  for (long long f = tid; f < nW; f += stride) {
    const int c = (int)((f / (Ch * 8 * WG)) % 2) * Ch + (int)(f % Ch);
    const int k4 = (int)((f / Ch) % 8);
    const int kc2 = (int)((f / (Cpad * 8 * WG)) % (4 / WG)) * WG + (int)((f / (Ch * 8)) % WG);
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

// Phase clocks (ADLM_BIG_EXP & 16): cycles each role of every CTA spent per phase, printed by the host after the launch.
constexpr int DBG_SLOTS = 24;
__device__ long long g_big_clocks[160][DBG_SLOTS];
template <bool DBG>
struct PhaseClock {
  long long t, acc[12];
  bool on;
  __device__ __forceinline__ void start(bool enable) {
    on = DBG && enable;
    for (int i = 0; i < 12; ++i) acc[i] = 0;
    t = on ? clock64() : 0;
  }
  __device__ __forceinline__ void tick(int i) {
    if (DBG && on) { const long long n = clock64(); acc[i] += n - t; t = n; }
  }
  __device__ __forceinline__ void flush(int base, int count) {
    if (DBG && on && (threadIdx.x & 31) == 0 && blockIdx.x < 160)
      for (int i = 0; i < count; ++i) g_big_clocks[blockIdx.x][base + i] = acc[i];
  }
};

// Event trace (ADLM_BIG_EXP & 64): CTA 0 logs (event, clock) of its MMA warp and of epilogue warp 0 during tiles 2-3.
constexpr int TRACE_N = 1024;
__device__ long long g_big_trace[2][TRACE_N];
__device__ int g_big_trace_ev[2][TRACE_N];
template <bool DBG>
struct Trace {
  int n, role;
  bool on;
  __device__ __forceinline__ void start(bool enable, int r) { on = DBG && enable; n = 0; role = r; }
  __device__ __forceinline__ void enable(bool e) { on = DBG && e; }
  __device__ __forceinline__ void ev(int id) {
    if (DBG) {
      if (on && n < TRACE_N && (threadIdx.x & 31) == 0) {
        g_big_trace[role][n] = clock64();
        g_big_trace_ev[role][n] = id;
      }
      if (on) ++n;
    }
  }
};

// barrier indices (same layout in both CTAs)
// The A tile is handed over in depth GROUPS (the same PG chunks a ring step covers; at most 4 groups): the MMAs of
// the tile's first prototype block start on group 0 while the producers still fill group 1, and a group is refilled
// for the next tile as soon as the LAST block's MMAs on it are done — under that block's remaining MMAs and the last
// two last-layer blocks — instead of after the whole tile (the MMA warp waited ~3.7 k cycles per tile for A).
constexpr int A_GROUPS_MAX = 4;
constexpr int B_AFULL = 0;               // [4] LEADER: depth group of both CTAs' A tiles written (4 + 4 producer warps)
constexpr int B_AFREE = 4;               // [4] the tile's last MMA1 on the group has read it (commit, both CTAs)
constexpr int B_D1FULL = 8;              // [2] D1[b] ready (commit, both CTAs)
constexpr int B_SREADY = 10;             // [2] LEADER: similarities back in D1[b] of both CTAs (8 + 8 epilogue warps)
constexpr int B_D2FULL = 12;             // [2] D2[tile parity] complete (commit, both CTAs)
constexpr int B_PFULL = 16;              // [P_SLOTS_MAX] own prototype chunk landed (expect_tx)
constexpr int B_PEMPTY = 32;             // [P_SLOTS_MAX] MMAs of the chunk done (commit, both CTAs)
constexpr int B_WFULL = 48;              // [W_SLOTS]
constexpr int B_WEMPTY = 52;             // [W_SLOTS]
constexpr int B_XPFULL = 56;             // [P_SLOTS_MAX] LEADER: the peer's chunk landed (peer's forwarder)
constexpr int B_XWFULL = 72;             // [W_SLOTS] LEADER: same for the last-layer chunks
constexpr int NBARS = 76;

__device__ __forceinline__ void bulk_load(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_smem), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

struct BigSmem {
  // A [KPAD/4][128][4] | P ring | W ring | p2 [Npad] | x2 [2][128] | barriers | tmem base
  static __host__ __device__ size_t a() { return 0; }
  static __host__ __device__ size_t pring(int KC) { return (size_t)KC * KCH * TILE * 4; }
  // (`slots` = ring slots x chunks per slot: the ring's size in 8 KB chunks)
  static __host__ __device__ size_t wring(int KC, int slots) { return pring(KC) + (size_t)slots * P_CHUNK_FLOATS * 4; }
  static __host__ __device__ size_t p2(int KC, int slots, int Cpad) { return wring(KC, slots) + (size_t)W_SLOTS * WG * 32 * (Cpad / 2) * 4; }
  static __host__ __device__ size_t x2(int KC, int slots, int Cpad, int Npad) { return p2(KC, slots, Cpad) + (size_t)Npad * 4; }
  static __host__ __device__ size_t bars(int KC, int slots, int Cpad, int Npad) { return x2(KC, slots, Cpad, Npad) + 2 * TILE * 4; }
  static __host__ __device__ size_t total(int KC, int slots, int Cpad, int Npad) { return bars(KC, slots, Cpad, Npad) + NBARS * 8 + 16; }
};

template <bool DBG, int PG>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(BIG_THREADS, 1) k_tc_fwd_big(const BigParams p) {
  constexpr int P_GROUP_FLOATS = PG * P_CHUNK_FLOATS;
  const int xf = DBG ? p.exp_flags : 0;                            // experiment flags exist in the debug build only
  extern __shared__ __align__(1024) uint8_t smem[];
  const int KC = p.KC, nblk = p.nblk, Cpad = p.Cpad, Npad = nblk * NBLK, K = p.K, N = p.N, C = p.C, HW = p.HW;
  const uint32_t P_SLOTS = (uint32_t)p.p_slots;
  float* A = reinterpret_cast<float*>(smem + BigSmem::a());
  float* Pr = reinterpret_cast<float*>(smem + BigSmem::pring(KC));
  float* Wr = reinterpret_cast<float*>(smem + BigSmem::wring(KC, p.p_slots * PG));
  float* p2s = reinterpret_cast<float*>(smem + BigSmem::p2(KC, p.p_slots * PG, Cpad));
  float* x2s = reinterpret_cast<float*>(smem + BigSmem::x2(KC, p.p_slots * PG, Cpad, Npad));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + BigSmem::bars(KC, p.p_slots * PG, Cpad, Npad));
  uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(bars + NBARS);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int Mi = (int)p.M;
  // pair q of the grid's pairs handles the tile pairs q, q + pairs, ...; this CTA takes tile 2 * pair + rank (a tile
  // beyond the map has no valid pixel: it computes on the centre value and stores nothing)
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pairs = (int)gridDim.x / 2, pair0 = (int)blockIdx.x / 2, num_pairs = (p.num_tiles + 1) / 2;
  const int iters = num_pairs > pair0 ? (num_pairs - 1 - pair0) / pairs + 1 : 0;
  auto tile_of = [&](int it) { return 2 * (pair0 + it * pairs) + (int)rank; };
  const int Ch = Cpad / 2;                                         // class rows of a W chunk held by one CTA
  const uint32_t w_chunk_bytes = (uint32_t)(32 * Ch * 4), w_group_bytes = WG * w_chunk_bytes;
  const int KG = KC / PG;                                          // prototype-chunk groups per block

  if (warp == 12) tmem_alloc_pair(smem_u32(tmem_base_s), 512);
  if (tid == 0) {
    for (int i = 0; i < A_GROUPS_MAX; ++i) {
      mbar_init(smem_u32(&bars[B_AFULL + i]), 8);
      mbar_init(smem_u32(&bars[B_AFREE + i]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&bars[B_D1FULL + i]), 1);
      mbar_init(smem_u32(&bars[B_SREADY + i]), 16);
      mbar_init(smem_u32(&bars[B_D2FULL + i]), 1);
    }
    for (int i = 0; i < P_SLOTS_MAX; ++i) {
      mbar_init(smem_u32(&bars[B_PFULL + i]), 1);
      mbar_init(smem_u32(&bars[B_PEMPTY + i]), 1);
      mbar_init(smem_u32(&bars[B_XPFULL + i]), 1);
    }
    for (int i = 0; i < W_SLOTS; ++i) {
      mbar_init(smem_u32(&bars[B_WFULL + i]), 1);
      mbar_init(smem_u32(&bars[B_WEMPTY + i]), 1);
      mbar_init(smem_u32(&bars[B_XWFULL + i]), 1);
    }
    mbar_init_fence();
  }
  for (int n = tid; n < Npad; n += BIG_THREADS) p2s[n] = __ldg(p.p2 + n);
  fence_before();
  cluster_sync_all();                      // barriers of BOTH CTAs initialised before anyone arrives remotely
  fence_after();
  const uint32_t tmem_base = *tmem_base_s;
  const uint32_t bar0 = smem_u32(&bars[0]);
  auto bar = [&](int idx) { return bar0 + (uint32_t)idx * 8u; };
  const uint32_t tD1 = tmem_base, tD2 = tmem_base + 2 * NBLK;     // D1[b] at b*128, D2[t&1] at 256 + (t&1)*Cpad

  if (warp >= 8 && warp < 12) {
    // ================= producers: thread = pixel =================
    const int cw = tid - 256;
    PhaseClock<DBG> pc;                         // 0 wait for A to be free 1 fill
    pc.start((xf & 16) != 0 && warp == 8);
    for (int it = 0; it < iters; ++it) {
      const int m = tile_of(it) * TILE + cw;
      const bool valid = m < Mi;
      const int b = valid ? m / HW : 0;
      const float* src = p.z + ((size_t)b * K) * HW + (valid ? m - b * HW : 0);
      // Two depth chunks are kept in flight in registers: the fill of a tile is a chain of load round trips (the map
      // streams from DRAM), and it is the bubble the MMA warp sees at every tile boundary.
      float zr[2][KCH];
      auto load = [&](int kc, float (&dst)[KCH]) {
#pragma unroll
        for (int k = 0; k < KCH; ++k) {
          const int kk = kc * KCH + k;
          dst[k] = (valid && kk < K) ? __ldg(src + (size_t)kk * HW) : CENTRE;
        }
      };
      float x2 = 0.f;
      auto put = [&](int kc, float (&cur)[KCH]) {
        float v[KCH];
#pragma unroll
        for (int k = 0; k < KCH; ++k) {
          v[k] = rna_tf32_mask(cur[k] - CENTRE);
          x2 = fmaf(v[k], v[k], x2);
        }
        if (kc + 2 < KC) load(kc + 2, cur);
#pragma unroll
        for (int k4 = 0; k4 < KCH / 4; ++k4)
          *reinterpret_cast<float4*>(A + ((size_t)(kc * (KCH / 4) + k4) * TILE + cw) * 4) =
              make_float4(v[k4 * 4], v[k4 * 4 + 1], v[k4 * 4 + 2], v[k4 * 4 + 3]);
      };
      load(0, zr[0]);                                            // in flight while the previous tile still owns A
      load(1, zr[1]);                                            // (KC is even: a multiple of PG)
      for (int kc = 0; kc < KC; kc += 2) {
        const int g = kc / PG;
        if (kc % PG == 0 && it > 0) mbar_wait(bar(B_AFREE + g), (uint32_t)(it - 1) & 1u);
        if (kc == 0) pc.tick(0);
        put(kc, zr[0]);
        put(kc + 1, zr[1]);
        if ((kc + 2) % PG == 0) {                                // the group is complete: hand it to the MMA warp
          if (kc + 2 == KC) x2s[(it & 1) * TILE + cw] = x2;      // |z|^2 goes with the last group
          fence_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(bar(B_AFULL + g), 0);
        }
      }
      pc.tick(1);
      // pull the NEXT tile's lines towards L2 now (a whole tile period ahead): its fill then pays L2, not DRAM, latency.
      // Lane l touches the warp's 128-byte line of depth l, l + 32, ...
      if (it + 1 < iters) {
        const int mn = tile_of(it + 1) * TILE + (warp - 8) * 32;
        if (mn < Mi) {
          const int bn = mn / HW;
          const float* nsrc = p.z + ((size_t)bn * K) * HW + (mn - bn * HW);
          for (int kk = lane; kk < K; kk += 32)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(nsrc + (size_t)kk * HW));
        }
      }
    }
    pc.flush(11, 2);
  } else if (warp == 13) {
    // ================= prototype-chunk loader (one thread, TMA bulk copies) =================
    if (lane == 0) {
      uint32_t i = 0;
      PhaseClock<DBG> pc;                       // 0 wait for a free slot 1 issue
      pc.start((xf & 16) != 0);
      for (int it = 0; it < iters; ++it)
        for (int blk = 0; blk < nblk; ++blk)
          for (int kg = 0; kg < KG; ++kg, ++i) {
            const uint32_t slot = i % P_SLOTS, use = i / P_SLOTS;
            if (use > 0) mbar_wait(bar(B_PEMPTY + slot), (use - 1) & 1u);
            pc.tick(0);
            bulk_load(smem_u32(Pr + (size_t)slot * P_GROUP_FLOATS),
                      p.Pst + (((size_t)blk * KG + kg) * 2 + rank) * P_GROUP_FLOATS, P_GROUP_FLOATS * 4, bar(B_PFULL + slot));
            pc.tick(1);
            if (xf & 32) {          // serialise: measures the latency of one bulk copy (issue -> complete_tx)
              mbar_wait(bar(B_PFULL + slot), use & 1u);
              pc.tick(2);
            }
          }
      pc.flush(13, 3);
    }
  } else if (warp == 14) {
    // ================= last-layer-chunk loader =================
    if (lane == 0) {
      uint32_t i = 0;
      for (int it = 0; it < iters; ++it)
        for (int blk = 0; blk < nblk; ++blk)
          for (int wg = 0; wg < 4 / WG; ++wg, ++i) {
            const uint32_t slot = i % W_SLOTS, use = i / W_SLOTS;
            if (use > 0) mbar_wait(bar(B_WEMPTY + slot), (use - 1) & 1u);
            bulk_load(smem_u32(Wr + (size_t)slot * WG * 32 * Ch),
                      p.Wst + (((size_t)blk * (4 / WG) + wg) * 2 + rank) * WG * 32 * Ch, w_group_bytes, bar(B_WFULL + slot));
          }
    }
  } else if (warp == 12 && leader) {
    // ================= MMA issue (leader CTA; whole warp, one elected lane per instruction) =================
    const uint32_t idesc1 = make_idesc_tf32(2 * TILE, NBLK), idesc2 = make_idesc_tf32(2 * TILE, Cpad);
    const uint32_t a_base = smem_u32(A), p_base = smem_u32(Pr), w_base = smem_u32(Wr);
    // Ring positions as (slot, parity) counters; `*_ready` = the NEXT group's barriers were already seen complete.
    // The look at the next group's barriers sits in the MIDDLE of the current group's MMAs: a barrier round trip costs
    // ~150 cycles and the tensor pipe queues only a few MMAs, so with the look between two groups the pipe ran dry
    // for ~250 cycles per group (122 instead of 87 cycles per MMA in the event trace, profiles/r2_big_trace.log).
    uint32_t p_slot = 0, p_par = 0, w_slot = 0, w_par = 0;
    uint32_t d1par[2] = {0u, 0u};
    bool p_ready = false, w_ready = false, s_ready = false;
    const bool no_peek = (xf & 256) != 0;
    PhaseClock<DBG> pc;                         // 0 AFULL 1 PFULL 2 XPFULL 3 MMA1 issue 4 SREADY 5 WFULL 6 XWFULL 7 MMA2 issue
    pc.start((xf & 16) != 0);
    Trace<DBG> tr;
    tr.start(false, 0);
    // (the vote makes the flag warp-uniform for the compiler: the MMAs that follow stay in convergent code)
    auto peek_p = [&](uint32_t slot, uint32_t par) {
      if (!no_peek)
        p_ready = __all_sync(0xffffffffu, mbar_test(bar(B_PFULL + slot), par) && mbar_test(bar(B_XPFULL + slot), par));
    };
    auto peek_w = [&](uint32_t slot, uint32_t par) {
      if (!no_peek)
        w_ready = __all_sync(0xffffffffu, mbar_test(bar(B_WFULL + slot), par) && mbar_test(bar(B_XWFULL + slot), par));
    };
    // descriptors: high words are constant per operand, low words step by (bytes >> 4) from MMA to MMA
    const uint32_t a_hi = (uint32_t)(make_desc(0, TILE * 16, 128) >> 32), a_lo0 = (uint32_t)make_desc(a_base, TILE * 16, 128);
    const uint32_t p_hi = (uint32_t)(make_desc(0, NHALF * 16, 128) >> 32), p_lo0 = (uint32_t)make_desc(p_base, NHALF * 16, 128);
    const uint32_t w_hi = (uint32_t)(make_desc(0, Ch * 16, 128) >> 32), w_lo0 = (uint32_t)make_desc(w_base, Ch * 16, 128);
    constexpr uint32_t A_STEP = (2 * TILE * 16) >> 4, P_STEP = (2 * NHALF * 16) >> 4;   // per k-step of 8
    const uint32_t w_step = (uint32_t)(2 * Ch * 16) >> 4;
    const bool skip1 = (xf & 1) != 0, skip2 = (xf & 2) != 0;
    // MMA1 of one prototype block.  `next_s` = D1 buffer whose SREADY the following mma2 will need (-1: none).
    auto mma1 = [&](int it, int blk, bool last_of_tile, int next_s) {
      const int b = blk & 1;
      for (int kg = 0; kg < KG; ++kg) {
        const uint32_t slot = p_slot, par = p_par;
        if (!p_ready) {
          mbar_wait(bar(B_PFULL + slot), par);
          pc.tick(1);
          tr.ev(100 + blk * 10 + kg);           // own chunk group present
          mbar_wait(bar(B_XPFULL + slot), par);
          pc.tick(2);
        }
        tr.ev(200 + blk * 10 + kg);             // both present
        if (blk == 0) {                         // first block of the tile: this depth group of A must have been written
          mbar_wait(bar(B_AFULL + kg), (uint32_t)it & 1u);
          pc.tick(0);
        }
        p_ready = false;
        if (++p_slot == P_SLOTS) { p_slot = 0; p_par ^= 1u; }
        fence_after();
#pragma unroll
        for (int j = 0; j < 4 * PG; ++j) {
          const int ks = kg * (4 * PG) + j;
          if (j == 2 * PG) {                    // half of the group is queued: look ahead
            peek_p(p_slot, p_par);
            if (kg == KG - 1 && next_s >= 0 && !no_peek) {
              s_ready = __all_sync(0xffffffffu, mbar_test(bar(B_SREADY + next_s), d1par[next_s]));
              // the look from the previous block's last-layer MMAs came too early for a two-slot ring (the slot had
              // just been freed): look again now that the reload had a whole MMA1 block to land
              if (!w_ready) peek_w(w_slot, w_par);
            }
          }
          if (skip1) continue;
          // a chunk group is contiguous in the slot: [PG][8 slabs][64 rows][4] -> one step per MMA
          mma_pair_ss_lh(tD1 + b * NBLK, a_lo0 + (uint32_t)ks * A_STEP,
                         p_lo0 + slot * ((P_GROUP_FLOATS * 4) >> 4) + (uint32_t)j * P_STEP, a_hi, p_hi, idesc1, ks > 0 ? 1u : 0u);
        }
        mma_pair_commit_w(bar(B_PEMPTY + slot));
        if (last_of_tile) mma_pair_commit_w(bar(B_AFREE + kg));
        pc.tick(3);
        tr.ev(300 + blk * 10 + kg);             // MMA1 group issued
      }
      mma_pair_commit_w(bar(B_D1FULL + b));
      pc.tick(3);
    };
    auto mma2 = [&](int it, int blk) {
      const int b = blk & 1, tp = it & 1;
      if (!s_ready) mbar_wait(bar(B_SREADY + b), d1par[b]);
      s_ready = false;
      d1par[b] ^= 1u;
      pc.tick(4);
      tr.ev(400 + blk * 10);                    // similarities ready
      fence_after();
      for (int wg = 0; wg < 4 / WG; ++wg) {
        const uint32_t slot = w_slot, par = w_par;
        if (!w_ready) {
          mbar_wait(bar(B_WFULL + slot), par);
          pc.tick(5);
          mbar_wait(bar(B_XWFULL + slot), par);
          pc.tick(6);
        }
        tr.ev(500 + blk * 10 + wg);             // W group present (both)
        w_ready = false;
        if (++w_slot == W_SLOTS) { w_slot = 0; w_par ^= 1u; }
        fence_after();
#pragma unroll
        for (int j = 0; j < 4 * WG; ++j) {
          if (j == 2 * WG) {
            peek_w(w_slot, w_par);
            if (wg == 4 / WG - 1) peek_p(p_slot, p_par);       // the next prototype group, whichever block it belongs to
          }
          if (!skip2)
            mma_pair_ts_lh(tD2 + tp * Cpad, tD1 + b * NBLK + (wg * (4 * WG) + j) * 8,
                           w_lo0 + slot * (w_group_bytes >> 4) + (uint32_t)j * w_step, w_hi, idesc2,
                           (blk > 0 || wg > 0 || j > 0) ? 1u : 0u);
        }
        mma_pair_commit_w(bar(B_WEMPTY + slot));
        pc.tick(7);
        tr.ev(600 + blk * 10 + wg);             // MMA2 group issued
      }
      if (blk == nblk - 1) mma_pair_commit_w(bar(B_D2FULL + tp));
    };
    for (int it = 0; it < iters; ++it) {
      tr.enable((xf & 64) != 0 && blockIdx.x == 0 && (it == 2 || it == 3));
      tr.ev(10);                                // tile start
      mma1(it, 0, nblk == 1, nblk == 1 ? 0 : -1);
      for (int blk = 0; blk < nblk; ++blk) {
        if (blk + 1 < nblk) mma1(it, blk + 1, blk + 1 == nblk - 1, blk & 1);
        mma2(it, blk);
      }
    }
    pc.flush(0, 8);
  } else if (warp == 12) {
    // ================= peer CTA: forwarder =================
    // Walks the chunk sequence in the leader's issue order; when one of this CTA's chunks has landed it arrives on the
    // leader's barrier for that slot.  (The leader cannot wait on a barrier in another CTA, and a bulk copy signals a
    // barrier of the CTA it writes to.)  A slot is reloaded only after the leader's commit for its previous use, which
    // follows the leader's wait on that use — the forwarder is never more than one phase ahead.
    if (lane == 0) {
      uint32_t pi = 0, wi = 0;
      auto fwd_p = [&]() {
        for (int kg = 0; kg < KG; ++kg, ++pi) {
          const uint32_t slot = pi % P_SLOTS, use = pi / P_SLOTS;
          mbar_wait(bar(B_PFULL + slot), use & 1u);
          mbar_arrive_cluster(bar(B_XPFULL + slot), 0);
        }
      };
      auto fwd_w = [&]() {
        for (int wg = 0; wg < 4 / WG; ++wg, ++wi) {
          const uint32_t slot = wi % W_SLOTS, use = wi / W_SLOTS;
          mbar_wait(bar(B_WFULL + slot), use & 1u);
          mbar_arrive_cluster(bar(B_XWFULL + slot), 0);
        }
      };
      for (int it = 0; it < iters; ++it) {
        fwd_p();
        for (int blk = 0; blk < nblk; ++blk) {
          if (blk + 1 < nblk) fwd_p();
          fwd_w();
        }
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
    PhaseClock<DBG> pc;                         // 0 D1FULL wait 1 similarity work + arrive 2 logits read-out
    pc.start((xf & 16) != 0 && warp == 0);
    Trace<DBG> tr;
    tr.start(false, 1);

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
        tr.enable((xf & 64) != 0 && blockIdx.x == 0 && warp == 0 && (it == 2 || it == 3));
        mbar_wait(bar(B_D1FULL + db), d1cnt[db] & 1u);
        ++d1cnt[db];
        pc.tick(0);
        tr.ev(700 + blk * 10);                  // D1 complete
        fence_after();
        if (blk == 0) x2 = x2s[(it & 1) * TILE + row];
        const uint32_t t1 = tD1 + db * NBLK + lane_field;
#pragma unroll 1
        for (int ch = 0; ch < 4; ++ch) {
          if (xf & 4) break;
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
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(bar(B_SREADY + db), 0);
        pc.tick(1);
        tr.ev(800 + blk * 10);                  // similarities written

        // logits of the previous tile (its D2 is complete since its last block): the two halves alternate
        if (blk == 0 && it > 0 && ((it - 1) & 1) == half) { epilogue2(it - 1); tr.ev(900); }
        pc.tick(2);
      }
    }
    if (iters > 0 && ((iters - 1) & 1) == half) epilogue2(iters - 1);
    pc.tick(2);
    pc.flush(8, 3);
  }
  fence_before();
  cluster_sync_all();                      // no CTA leaves while its peer may still arrive on, or read from, it
  if (warp == 12) tmem_dealloc_pair(tmem_base, 512);
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
  const int PG = pick_pg(K);
  const int KC = (K + KCH * PG - 1) / (KCH * PG) * PG, nblk = (N + NBLK - 1) / NBLK, Cpad = (C + 15) & ~15;
  return align_up((size_t)nblk * KC * 2 * P_CHUNK_FLOATS * 4, 256) + align_up((size_t)nblk * NBLK * 4, 256) +
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
  const int PG = pick_pg(s.K);
  prm.PG = PG;
  prm.KC = (s.K + KCH * PG - 1) / (KCH * PG) * PG;
  prm.nblk = (s.N + NBLK - 1) / NBLK;
  prm.Cpad = (s.C + 15) & ~15;
  char* w8 = static_cast<char*>(workspace);
  float* Pst = reinterpret_cast<float*>(w8);
  float* p2 = reinterpret_cast<float*>(w8 + align_up((size_t)prm.nblk * prm.KC * 2 * P_CHUNK_FLOATS * 4, 256));
  float* Wst = reinterpret_cast<float*>(reinterpret_cast<char*>(p2) + align_up((size_t)prm.nblk * NBLK * 4, 256));
  prm.z = z; prm.Pst = Pst; prm.p2 = p2; prm.Wst = Wst;
  prm.logits = logits; prm.dist = dist; prm.argmax = argmax;
  prm.B = s.B; prm.K = s.K; prm.HW = s.HW; prm.N = s.N; prm.C = s.C;
  prm.eps = s.eps; prm.act_mode = s.act_mode;
  prm.M = (long long)s.B * s.HW;
  prm.num_tiles = (int)((prm.M + TILE - 1) / TILE);
  prm.exp_flags = 0;
  if (const char* e = getenv("ADLM_BIG_EXP")) prm.exp_flags = atoi(e);
  const size_t fixed = BigSmem::total(prm.KC, 0, prm.Cpad, prm.nblk * NBLK);
  const size_t room = fixed < 227 * 1024 ? 227 * 1024 - fixed : 0;
  prm.p_slots = (int)(room / ((size_t)PG * P_CHUNK_FLOATS * 4));
  if (prm.p_slots > P_SLOTS_MAX) prm.p_slots = P_SLOTS_MAX;
  if (const char* e = getenv("ADLM_BIG_SLOTS")) { const int v = atoi(e); if (v >= 2 && v <= prm.p_slots) prm.p_slots = v; }
  if (prm.p_slots < 2) {
    set_error("adlm_head_fwd_big: %zu bytes of shared memory needed before the prototype ring (> 227 KB - 16 KB)", fixed);
    return ADLM_E_UNSUPPORTED;
  }
  const size_t bytes = BigSmem::total(prm.KC, prm.p_slots * PG, prm.Cpad, prm.nblk * NBLK);
  k_big_stage<<<256, 256, 0, st>>>(P, W, s.N, s.K, s.C, prm.KC, PG, prm.nblk, prm.Cpad,
                                   s.act_mode == ADLM_ACT_LOG ? LN2 : 1.f, Pst, p2, Wst);
  ADLM_LAUNCH_CHECK();
  auto kernel = prm.exp_flags != 0 ? (PG == 4 ? k_tc_fwd_big<true, 4> : k_tc_fwd_big<true, 2>)
                                   : (PG == 4 ? k_tc_fwd_big<false, 4> : k_tc_fwd_big<false, 2>);
  ADLM_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  // one CTA pair per TPC (the kernel carries __cluster_dims__(2, 1, 1)): an even grid of at most the SM count
  const int num_pairs = (prm.num_tiles + 1) / 2, max_pairs = sm_count_big() / 2;
  const int grid = 2 * (num_pairs < max_pairs ? num_pairs : max_pairs);
  KernelTimer timer(ADLM_K_HEAD_FWD, st);
  kernel<<<grid, BIG_THREADS, bytes, st>>>(prm);
  ADLM_LAUNCH_CHECK();
  if (prm.exp_flags & 64) {
    static long long ht[2][TRACE_N];
    static int he[2][TRACE_N];
    cudaStreamSynchronize(st);
    cudaMemcpyFromSymbol(ht, g_big_trace, sizeof(ht));
    cudaMemcpyFromSymbol(he, g_big_trace_ev, sizeof(he));
    long long t0 = 0;
    for (int i = 0; i < TRACE_N; ++i) if (he[0][i] == 10) { t0 = ht[0][i]; break; }
    for (int r = 0; r < 2; ++r)
      for (int i = 0; i < TRACE_N; ++i)
        if (he[r][i] != 0) fprintf(stderr, "[big trace] %s %d %lld\n", r ? "epi" : "mma", he[r][i], ht[r][i] - t0);
    cudaMemset(nullptr, 0, 0);
  }
  if (prm.exp_flags & 16) {
    static long long h[160][DBG_SLOTS];
    cudaStreamSynchronize(st);
    cudaMemcpyFromSymbol(h, g_big_clocks, sizeof(h));
    static const char* names[16] = {"mma:AFULL", "mma:PFULL", "mma:XPFULL", "mma:MMA1issue", "mma:SREADY", "mma:WFULL",
                                    "mma:XWFULL", "mma:MMA2issue", "epi:D1FULL", "epi:work", "epi:logits", "prod:AFREE",
                                    "prod:fill", "pload:PEMPTY", "pload:issue", "pload:latency(exp32)"};
    for (int c = 0; c < 4 && c < grid; ++c) {
      fprintf(stderr, "[big clocks] cta %d:", c);
      for (int i = 0; i < 16; ++i) fprintf(stderr, " %s=%lld", names[i], h[c][i]);
      fprintf(stderr, "\n");
    }
  }
  return ADLM_OK;
}

}  // namespace adlm
