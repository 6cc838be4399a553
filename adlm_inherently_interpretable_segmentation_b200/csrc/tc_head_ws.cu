// Prototype-head forward, ADLM_PREC_TF32, as a warp-specialised tcgen05 pipeline (model.py:203-283).
//
// Per 128-pixel tile (TMEM lanes = pixels, columns = prototypes):
//   producer warps 8-11   thread = pixel: LDG z[k][px] one tile ahead (NCHW, coalesced), centre by 1/2, round to
//                         nearest TF32, write the K-major A operand tile (two stages) and |z|^2
//   MMA warp 12           MMA1  D1[a] = A . Pk^T           (SS; accumulator stage a = tile & 1)
//                         MMA2  D2[a] = S . Wk^T           (.ts: S read from TMEM where epilogue 1 left it)
//   epilogue warps 0-7    lane quarter = warp % 4, column half = warp / 4
//                         E1: tcgen05.ld D1 -> d = max(|z|^2 - 2 acc + |p|^2, 0) -> s = log2((d+1)/(d+eps)) ->
//                             tcgen05.st over D1 (+ optional d / s / own-class-distance stores)
//                         E2 (one of the two halves per tile, alternating): logits from D2 -> staged through
//                             shared memory -> 16-byte coalesced stores, per-pixel argmax
// MMA1 of tile i+1 and MMA2 of tile i-1 run while the eight epilogue warps work on tile i, and the producers
// are a tile ahead: in steady state the kernel is bound by the epilogue's two MUFU.LG2 per (pixel, prototype).
// ln 2 is folded into Wk, so the similarity leaves the epilogue in log2 units.
#include "tc_head.cuh"
#include "tc_common.cuh"

namespace adlm {

namespace {

using namespace tc;

constexpr int TILE = 128;
constexpr int JMAX = ADLM_MAX_PROTOS_PER_CLASS;
constexpr float LN2 = 0.6931471805599453f;
constexpr float CENTRE = 0.5f;
constexpr int WS_THREADS = 416;   // 8 epilogue + 4 producer + 1 MMA warp

// -DADLM_PHASE_CLOCKS: CTA 0 records clock64() at phase boundaries (roles: 0 epilogue half 0, 1 epilogue half 1,
// 2 producer, 3 MMA issuer); read back with adlm_debug_ws_phase_clocks (tools/phase_clocks_ws.py)
#ifdef ADLM_PHASE_CLOCKS
__device__ long long g_ws_phase[4][64];
#define WS_PHASE(role, i)                                                                   \
  do {                                                                                      \
    if (blockIdx.x == 0 && lane == 0 && (i) < 64) g_ws_phase[(role)][(i)] = clock64();       \
  } while (0)
#else
#define WS_PHASE(role, i)
#endif

// barrier indices ([2] = per accumulator / operand stage)
constexpr int B_FULL = 0;      // A operand stage written (128 producer threads)
constexpr int B_SFREE = 2;     // MMA1 done reading the operand stage (commit)
constexpr int B_D1FULL = 4;    // D1 ready (commit)
constexpr int B_SREADY = 6;    // similarities written back to TMEM (256 epilogue threads)
constexpr int B_D2FULL = 8;    // D2 ready (commit)
constexpr int B_XFREE = 10;    // |z|^2 slot read by every epilogue thread (256)
constexpr int B_META = 12;     // |z|^2 slot written (128 producer threads)
constexpr int NBARS = 14;

template <int KP>
struct WsSmem {
  static __host__ __device__ size_t pk() { return 0; }
  static __host__ __device__ size_t wk(int Npad) { return (size_t)KP * Npad * 4; }
  static __host__ __device__ size_t abuf(int stage, int Npad, int Cpad) {
    return wk(Npad) + (size_t)Npad * Cpad * 4 + (size_t)stage * KP * TILE * 4;
  }
  static __host__ __device__ size_t stg(int half, int Npad, int Cpad) { return abuf(2, Npad, Cpad) + (size_t)half * TILE * 32 * 4; }
  static __host__ __device__ size_t misc(int Npad, int Cpad) { return stg(2, Npad, Cpad); }
  // p2 [Npad] f32, class [Npad] i32, rank [Npad] i32, chunk class masks [16] u32, x2 [2][128] f32, barriers, tmem base
  static __host__ __device__ size_t total(int Npad, int Cpad) { return misc(Npad, Cpad) + (size_t)Npad * 12 + 64 + 1024 + NBARS * 8 + 16; }
};

template <int KP>
// 13 warps: the sub-partition that hosts four of them caps the kernel at 16384 / (4 * 32) = 128 registers
__global__ void __launch_bounds__(WS_THREADS, 1) k_tc_fwd_ws(const TcFwdParams p) {
  using L = WsSmem<KP>;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int Npad = p.Npad, Cpad = p.Cpad, K = p.K, N = p.N, C = p.C, HW = p.HW;
  float* Pk = reinterpret_cast<float*>(smem + L::pk());
  float* Wk = reinterpret_cast<float*>(smem + L::wk(Npad));
  uint8_t* misc = smem + L::misc(Npad, Cpad);
  float* p2s = reinterpret_cast<float*>(misc);
  int* pcs = reinterpret_cast<int*>(misc + (size_t)Npad * 4);
  int* prs = reinterpret_cast<int*>(misc + (size_t)Npad * 8);
  uint32_t* cmask = reinterpret_cast<uint32_t*>(misc + (size_t)Npad * 12);
  float* x2s = reinterpret_cast<float*>(misc + (size_t)Npad * 12 + 64);          // [2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(misc + (size_t)Npad * 12 + 64 + 1024);
  uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(bars + NBARS);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool want_own = p.own_dist != nullptr;
  const int Mi = (int)p.M;
  const int iters = p.num_tiles > (int)blockIdx.x ? (p.num_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  auto tile_of = [&](int it) { return (int)blockIdx.x + it * (int)gridDim.x; };

  if (warp == 12) tmem_alloc(smem_u32(tmem_base_s), 512);
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&bars[B_FULL + i]), 128);
      mbar_init(smem_u32(&bars[B_SFREE + i]), 1);
      mbar_init(smem_u32(&bars[B_D1FULL + i]), 1);
      mbar_init(smem_u32(&bars[B_SREADY + i]), 256);
      mbar_init(smem_u32(&bars[B_D2FULL + i]), 1);
      mbar_init(smem_u32(&bars[B_XFREE + i]), 256);
      mbar_init(smem_u32(&bars[B_META + i]), 128);
    }
    mbar_init_fence();
  }
  const uint32_t bar0 = smem_u32(&bars[0]);
  auto bar = [&](int idx) { return bar0 + (uint32_t)idx * 8u; };

  // ---- producers fetch their first tile before the operands are staged ----
  const bool is_producer = warp >= 8 && warp < 12;
  const int cw = tid - 256;
  float zr[KP];
  auto prefetch = [&](int it) {
    const int m = tile_of(it) * TILE + cw;
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
    } else {
#pragma unroll
      for (int k = 0; k < KP; ++k) zr[k] = CENTRE;
    }
  };
  if (is_producer) prefetch(0);

  // ---- one-time operand staging (all 13 warps) ----
  stage_rows_kmajor<KP, false>(Pk, Pk, p.P, N, K, Npad, CENTRE, tid, WS_THREADS);
  // Wk: B operand of MMA2, rows = classes, K index = prototype: [Npad/4][Cpad][4]; ln 2 folded in
  stage_w_kmajor(Wk, p.W, C, N, Cpad, Npad, p.act_mode == ADLM_ACT_LOG ? LN2 : 1.f, tid, WS_THREADS);
  __syncthreads();
  // Everything up to the first logits epilogue is independent of the step's prologue kernel (k_step_prepare), under
  // which this grid may have started: TMEM, barriers, feature tiles, P and W, the first tiles' MMAs and similarity
  // epilogues.  The prototype ranks are recomputed here from proto_class (an input) instead of being read from the
  // prologue's output, so the wait for that grid sits where its outputs are first touched: the CE normaliser and the
  // cleared accumulators, in the logits epilogue (`waited` in epilogue2).
  for (int n = tid; n < Npad; n += WS_THREADS) {
    // |p|^2 of the ROUNDED vector, so that z == p gives d == 0
    float a = 0.f;
    for (int k4 = 0; k4 < KP / 4; ++k4) {
      const float4 v = *reinterpret_cast<const float4*>(Pk + ((size_t)k4 * Npad + n) * 4);
      a = fmaf(v.x, v.x, a); a = fmaf(v.y, v.y, a); a = fmaf(v.z, v.z, a); a = fmaf(v.w, v.w, a);
    }
    p2s[n] = a;
    // rank of the prototype inside its class = earlier prototypes of the same class (k_step_prepare's rule, loss.cu);
    // beyond the per-class capacity of the KLD side buffers it is -1 and the prologue kernel fails the step
    int rk = -1;
    if (want_own && n < N) {
      const int c = __ldg(p.proto_class + n);
      int r = 0;
      for (int m = 0; m < n; ++m) r += (__ldg(p.proto_class + m) == c);
      rk = r < JMAX ? r : -1;
    }
    pcs[n] = rk >= 0 ? p.proto_class[n] : -1;
    prs[n] = rk >= 0 ? rk : 0;
  }
  __syncthreads();
  for (int c = tid; c < Npad / 16; c += WS_THREADS) {
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
  const uint32_t tmem_base = *tmem_base_s;
  if (warp == 0 || warp == 4) WS_PHASE(warp >> 2, 0);
  if (warp == 8) WS_PHASE(2, 0);
  if (warp == 12) WS_PHASE(3, 0);
  // TMEM columns: D1[a] at a*Npad (a = tile & 1), D2[a] at 2*Npad + a*Cpad; 2*(Npad + Cpad) <= 512.
  // No "accumulator empty" barrier is needed: the tensor pipe executes the MMA warp's instructions in issue
  // order (MMA1(t+2) after MMA2(t), which was issued after all epilogue threads signalled S(t) ready), and the
  // epilogue group that reads D2(t) does so, in program order, before it signals S(t+2) ready.
  const uint32_t tD1 = tmem_base, tD2 = tmem_base + 2 * Npad;

  if (is_producer) {
    // ================= producer: thread = pixel =================
    for (int it = 0; it < iters; ++it) {
      const int a = it & 1, j = it >> 1;
      float* A = reinterpret_cast<float*>(smem + L::abuf(a, Npad, Cpad));
      if (warp == 8) WS_PHASE(2, 1 + it * 4);
      if (j > 0) mbar_wait(bar(B_SFREE + a), (uint32_t)(j - 1) & 1u);    // MMA1 of tile it-2 has read the stage
      if (warp == 8) WS_PHASE(2, 2 + it * 4);
      if (p.z_out != nullptr) {
        // fused add-on activation: the tile holds pre-sigmoid features; z is needed again by the backward
        // kernels and by the callers (forward_with_features, push), so it is written out once from here
        const int m = tile_of(it) * TILE + cw;
        const bool ok = m < Mi;
        const int bi = ok ? m / HW : 0;
        float* zo = p.z_out + ((size_t)bi * K) * HW + (m - bi * HW);
#pragma unroll
        for (int k = 0; k < KP; ++k) {
          const float zv = rcp_approx(1.f + ex2_approx(-1.4426950408889634f * zr[k]));
          if (ok && k < K) zo[(size_t)k * HW] = zv;
          zr[k] = (ok && k < K) ? zv : CENTRE;
        }
      }
      float x2 = 0.f;
#pragma unroll
      for (int k4 = 0; k4 < KP / 4; ++k4) {
        float v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          v[e] = rna_tf32_mask(zr[k4 * 4 + e] - CENTRE);
          x2 = fmaf(v[e], v[e], x2);
        }
        *reinterpret_cast<float4*>(A + ((size_t)k4 * TILE + cw) * 4) = make_float4(v[0], v[1], v[2], v[3]);
      }
      fence_async_smem();
      mbar_arrive(bar(B_FULL + a));
      if (j > 0) mbar_wait(bar(B_XFREE + a), (uint32_t)(j - 1) & 1u);    // |z|^2 of tile it-2 read by all
      x2s[a * 128 + cw] = x2;
      mbar_arrive(bar(B_META + a));
      if (warp == 8) WS_PHASE(2, 3 + it * 4);
      prefetch(it + 1);
      if (warp == 8) WS_PHASE(2, 4 + it * 4);
    }
  } else if (warp == 12) {
    // ================= MMA issuer (whole warp, one elected lane per instruction; one call site per MMA group) ======
    {
      const uint32_t idesc1 = make_idesc_tf32(TILE, Npad), idesc2 = make_idesc_tf32(TILE, Cpad);
      const int ksteps = (K + 7) >> 3;
      const uint64_t db_p = make_desc(smem_u32(Pk), Npad * 16, 128);
      const uint64_t db_w = make_desc(smem_u32(Wk), Cpad * 16, 128);
      const uint32_t a_kstep = (2u * TILE * 16u) >> 4, p_kstep = (uint32_t)(2 * Npad * 16) >> 4,
                     w_kstep = (uint32_t)(2 * Cpad * 16) >> 4;
#pragma unroll 1
      for (int it = 0; it <= iters; ++it) {
        if (it < iters) {
          const int a = it & 1, j = it >> 1;
          const uint64_t da = make_desc(smem_u32(smem + L::abuf(a, Npad, Cpad)), TILE * 16, 128);
          if (lane == 0) WS_PHASE(3, 1 + it * 4);
          mbar_wait(bar(B_FULL + a), (uint32_t)j & 1u);
          fence_after();
          if (lane == 0) WS_PHASE(3, 2 + it * 4);
          const uint32_t d1 = tD1 + a * Npad;
#pragma unroll 1
          for (int ks = 0; ks < ksteps; ++ks)
            mma_ss_w(d1, da + (uint64_t)(ks * a_kstep), db_p + (uint64_t)(ks * p_kstep), idesc1, ks > 0 ? 1u : 0u);
          mma_commit_w(bar(B_SFREE + a));
          mma_commit_w(bar(B_D1FULL + a));
          if (lane == 0) WS_PHASE(3, 3 + it * 4);
        }
        if (it > 0) {                            // last layer of the previous tile
          const int t = it - 1, b = t & 1, jb = t >> 1;
          const uint32_t d1 = tD1 + b * Npad, d2 = tD2 + b * Cpad;
          mbar_wait(bar(B_SREADY + b), (uint32_t)jb & 1u);
          fence_after();
#pragma unroll 1
          for (int ks = 0; ks < Npad / 8; ++ks)
            mma_ts_w(d2, d1 + ks * 8, db_w + (uint64_t)(ks * w_kstep), idesc2, ks > 0 ? 1u : 0u);
          mma_commit_w(bar(B_D2FULL + b));
        }
        if (lane == 0 && it < iters) WS_PHASE(3, 4 + it * 4);
      }
    }
  } else {
    // ================= epilogue warps =================
    const int q = warp & 3, half = warp >> 2;
    const int row = q * 32 + lane;
    const uint32_t lane_field = (uint32_t)(q * 32) << 16;
    const int nch = Npad / 16;
    const int ch_begin = half == 0 ? 0 : (nch + 1) / 2, ch_end = half == 0 ? (nch + 1) / 2 : nch;
    const bool act_log = p.act_mode == ADLM_ACT_LOG;
    const float eps = p.eps, one_m_eps = 1.f - p.eps;
    float* stg = reinterpret_cast<float*>(smem + L::stg(half, Npad, Cpad));

    // Data-parallel step: the loss normaliser is the sum over all ranks.  k_step_prepare has published this rank's
    // counts to every peer without waiting (DpComm::defer_counts); the peers' arrive in this rank's exchange buffer
    // while the first tiles' MMAs and similarity epilogues run, and are collected here, at first use: one lane per
    // warp acquires the `world` flags (normally already set), sums the slots in rank order — every warp of every rank
    // gets the same bits — and CTA 0 stores the sums for the kernels that follow the forward.
    bool have_counts = false, waited = false;
    int n_ce_global = 1;
    auto collect_counts = [&]() {
      // lane r looks at rank r's flag and slot (a system-scope acquire load costs ~1 us: never chain them)
      const int par = (int)(p.ce_dp.seq & 1u), Wd = p.ce_dp.world;
      int v = 0, t2 = 0, ok = 1;
      if (lane < Wd) {
        char* me = p.ce_dp.peer[p.ce_dp.rank];
        if (!dp_wait_flag(reinterpret_cast<const unsigned int*>(me + p.ce_dp_lay.cflag) + par * Wd + lane, p.ce_dp.seq,
                          p.ce_dp.timeout_ns))
          ok = 0;
        const volatile int* slot_in = reinterpret_cast<const volatile int*>(me + p.ce_dp_lay.cnt) + (par * Wd + lane) * 2;
        v = slot_in[0];
        t2 = slot_in[1];
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {          // integer sums: the order does not matter, every warp gets the same bits
        v += __shfl_xor_sync(0xffffffffu, v, o);
        t2 += __shfl_xor_sync(0xffffffffu, t2, o);
      }
      ok = __all_sync(0xffffffffu, ok);
      if (blockIdx.x == 0 && warp == 0 && lane == 0) {
        p.ce_counts_out[0] = v;
        p.ce_counts_out[1] = t2;
        if (!ok) p.ce_accum[3] = 1.0;       // a peer never showed up: the step is flagged as failed (scalars[7])
      }
      n_ce_global = v;
      have_counts = true;
    };

    auto epilogue2 = [&](int t) {
      const int b = t & 1, jb = t >> 1;
      const int tile = tile_of(t);
      const int m = tile * TILE + row;
      // label and CE normaliser travel while the last-layer MMA finishes
      int tg = -1;
      float ce_sc = 0.f;
      if (!waited) {                       // no-op unless launched with programmatic stream serialisation
        asm volatile("griddepcontrol.wait;" ::: "memory");
        waited = true;
      }
      if (p.ce_accum != nullptr) {
        if (m < Mi) tg = p.target[m];
        if (p.ce_dp.seq != 0u) {
          if (!have_counts) collect_counts();
          ce_sc = p.ce_scale / (float)n_ce_global;
        } else {
          ce_sc = p.ce_scale / (float)p.ce_counts[0];
        }
      }
      mbar_wait(bar(B_D2FULL + b), (uint32_t)jb & 1u);
      fence_after();
      const uint32_t d2 = tD2 + b * Cpad + lane_field;
      float lg[32];
      uint32_t r[16];
      tmem_ld16(d2, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 16; ++i) lg[i] = __uint_as_float(r[i]);
      if (Cpad > 16) {
        tmem_ld16(d2 + 16, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) lg[16 + i] = __uint_as_float(r[i]);
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) lg[16 + i] = 0.f;
      }
      int best = 0;
      float bv = lg[0];
#pragma unroll
      for (int c = 1; c < 32; ++c)
        if (c < C && lg[c] > bv) { bv = lg[c]; best = c; }
      if (m < Mi && p.argmax != nullptr) p.argmax[m] = best;
      const int m0 = tile * TILE;
      const int npx = min(TILE, Mi - m0);
      const int cnt = npx * C;
      const int t128 = tid & 127;
      // a [128 px][C] tile is contiguous in HBM: transpose through shared memory, store 16 bytes per thread
      auto store_tile = [&](float* base) {
#pragma unroll
        for (int c = 0; c < 32; ++c)
          if (c < C) stg[row * C + c] = lg[c];
        named_bar_sync(1 + half, 128);
        float* dst = base + (size_t)m0 * C;
        if ((cnt & 3) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {
          for (int i = t128; i < (cnt >> 2); i += 128)
            reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(stg)[i];
        } else {
          for (int i = t128; i < cnt; i += 128) dst[i] = stg[i];
        }
        named_bar_sync(1 + half, 128);         // the staging buffer is reused
      };
      if (p.logits != nullptr) store_tile(p.logits);
      if (p.ce_accum != nullptr) {
        // cross entropy over the non-void patches (module.py:156-165), accuracy (:210-211), and
        // dlogits = w_ce * grad_scale / n_patches * (softmax - onehot), all from the registers
        bool ok;
        if (p.ignore_void) { ok = tg >= 1 && tg <= C; tg -= 1; }
        else               { ok = tg >= 0 && tg < C; }
        float loss = 0.f, hit = 0.f;
        if (ok) {
          float se = 0.f, lt = 0.f;
#pragma unroll
          for (int c = 0; c < 32; ++c)
            if (c < C) {
              se += __expf(lg[c] - bv);
              if (c == tg) lt = lg[c];
            }
          const float lse = bv + __logf(se);
          loss = lse - lt;
          hit = (best == tg) ? 1.f : 0.f;
          const float sc = ce_sc;
#pragma unroll
          for (int c = 0; c < 32; ++c)
            if (c < C) lg[c] = sc * (__expf(lg[c] - lse) - (c == tg ? 1.f : 0.f));
        } else {
#pragma unroll
          for (int c = 0; c < 32; ++c) lg[c] = 0.f;
        }
        if (p.dlogits != nullptr) store_tile(p.dlogits);
        loss = warp_sum(loss);
        hit = warp_sum(hit);
        if (lane == 0) {
          if (loss != 0.f) atomicAdd(&p.ce_accum[0], (double)loss);
          if (hit != 0.f) atomicAdd(&p.ce_accum[1], (double)hit);
        }
      }
    };

    // One more trip than there are tiles: trip `it` turns the distances of tile it into similarities (it < iters) and
    // then takes the logits of tile it-1.  A single call site keeps ONE copy of the logits epilogue in the kernel — the
    // three warp roles share the instruction cache, and ncu attributes a fifth of the warps' time to instruction fetch.
#pragma unroll 1
    for (int it = 0; it <= iters; ++it) {
      if (it < iters) {
      const int a = it & 1, j = it >> 1;
      const int m = tile_of(it) * TILE + row;
      const bool valid = m < Mi;
      const int b = valid ? m / HW : 0;
      const int pix = valid ? m - b * HW : 0;
      int lab = -1;
      if (want_own && valid) lab = p.target[m];
      const bool lab_ok = lab >= 0 && lab < 32;

      if (q == 0) WS_PHASE(half, 1 + it * 4);
      mbar_wait(bar(B_META + a), (uint32_t)j & 1u);
      mbar_wait(bar(B_D1FULL + a), (uint32_t)j & 1u);
      fence_after();
      if (q == 0) WS_PHASE(half, 2 + it * 4);
      const float x2 = x2s[a * 128 + row];
      mbar_arrive(bar(B_XFREE + a));
      const uint32_t t1 = tD1 + a * Npad + lane_field;

      auto chunk = [&](int ch, const uint32_t (&r)[16]) {
        float xp[16];
#pragma unroll
        for (int v4 = 0; v4 < 4; ++v4) {
          const float4 pq = *reinterpret_cast<const float4*>(p2s + ch * 16 + v4 * 4);
          xp[v4 * 4] = x2 + pq.x; xp[v4 * 4 + 1] = x2 + pq.y; xp[v4 * 4 + 2] = x2 + pq.z; xp[v4 * 4 + 3] = x2 + pq.w;
        }
        float dv[16], sv[16];
        uint32_t sv_bits[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float d = fmaxf(fmaf(-2.f, __uint_as_float(r[i]), xp[i]), 0.f);
          dv[i] = d;
          // log2((d+1)/(d+eps)) = log2(1 + (1-eps)/(d+eps)): one FADD, one FFMA, MUFU.RCP + MUFU.LG2
          const float s = act_log ? lg2_approx(fmaf(one_m_eps, rcp_approx(d + eps), 1.f)) : -d;
          sv[i] = s;
          sv_bits[i] = __float_as_uint(rna_tf32_fast(s));
        }
        tmem_st16(t1 + ch * 16, sv_bits);
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
              if (i < nmax) dst[i] = act_log ? LN2 * sv[i] : sv[i];
          }
          if (lab_ok && ((cmask[ch] >> lab) & 1u)) {
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (pcs[ch * 16 + i] == lab) p.own_dist[(size_t)m * JMAX + prs[ch * 16 + i]] = dv[i];
          }
        }
      };
      {
        // TMEM loads run one chunk ahead of the arithmetic; the loop stays rolled (one copy of the chunk body:
        // the three warp roles share a small instruction cache)
        uint32_t r[16], rn[16];
        tmem_ld16(t1 + ch_begin * 16, r);
#pragma unroll 1
        for (int ch = ch_begin; ch < ch_end; ++ch) {
          tmem_ld_wait();
          if (ch + 1 < ch_end) tmem_ld16(t1 + (ch + 1) * 16, rn);
          chunk(ch, r);
#pragma unroll
          for (int i = 0; i < 16; ++i) r[i] = rn[i];
        }
      }
      tmem_st_wait();
      fence_before();
      mbar_arrive(bar(B_SREADY + a));
      if (q == 0) WS_PHASE(half, 3 + it * 4);
      }
      // logits of the previous tile: the two halves alternate
      if (it > 0 && ((it - 1) & 1) == half) epilogue2(it - 1);
      if (q == 0) WS_PHASE(half, 4 + it * 4);
    }
  }
  fence_before();
  __syncthreads();
  if (warp == 12) tmem_dealloc(tmem_base, 512);
}

int g_sms_ws = 0;
int sm_count_ws() {
  if (g_sms_ws == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sms_ws, cudaDevAttrMultiProcessorCount, dev);
    if (g_sms_ws <= 0) g_sms_ws = 148;
  }
  return g_sms_ws;
}

template <int KP>
int launch_ws(const TcFwdParams& prm, cudaStream_t st) {
  using L = WsSmem<KP>;
  const size_t bytes = L::total(prm.Npad, prm.Cpad);
  if (bytes > 227 * 1024) {
    set_error("tc_head_fwd: %zu bytes of shared memory needed (> 227 KB) for K=%d N=%d", bytes, prm.K, prm.N);
    return ADLM_E_UNSUPPORTED;
  }
  ADLM_CUDA(cudaFuncSetAttribute(k_tc_fwd_ws<KP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  const int grid = prm.num_tiles < sm_count_ws() ? prm.num_tiles : sm_count_ws();
  KernelTimer timer(ADLM_K_HEAD_FWD, st);
  if (prm.after_prepare) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(WS_THREADS);
    cfg.dynamicSmemBytes = bytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    ADLM_CUDA(cudaLaunchKernelEx(&cfg, k_tc_fwd_ws<KP>, prm));
  } else {
    k_tc_fwd_ws<KP><<<grid, WS_THREADS, bytes, st>>>(prm);
  }
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

}  // namespace

#ifdef ADLM_PHASE_CLOCKS
extern "C" ADLM_API int adlm_debug_ws_phase_clocks(long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, g_ws_phase, sizeof(long long) * 256) == cudaSuccess ? 0 : -3;
}
#endif

int tc_head_fwd_ws(const TcFwdParams& prm, cudaStream_t st) {
  if (prm.K <= 32) return launch_ws<32>(prm, st);
  return launch_ws<64>(prm, st);
}

}  // namespace adlm
