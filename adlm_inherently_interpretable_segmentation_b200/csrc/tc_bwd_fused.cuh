// Fused ADLM_PREC_TF32 backward: ONE kernel forms G (and S) once per (pixel, prototype) and feeds all three
// contractions from it.  Included by tc_bwd.cu inside its anonymous namespace.
//
//   lanes = prototypes (the transposed problem D^T = Pc Zc^T), 64-pixel tiles, units = (tile, 128-prototype block)
//     D^T, Gl^T          tcgen05.mma SS into one of two TMEM buffers            (Gl = dlogits W)
//     epilogue           G = Gl ds/dd(d) [+ KLD own-class gradients] masked by relu, S = sim(d); both go back into the
//                        TMEM buffer (A operands) and G also into shared memory as a K-major operand [pixel][prototype]
//     dP += G Z          .ts MMA, persistent TMEM accumulator [prototype][depth]     (contraction over pixels)
//     dW += S dL         .ts MMA, persistent TMEM accumulator [prototype][class]
//     dZ' = G [Pc | 1]     SS MMA, A = the shared-memory G (rows = pixels), B = transposed prototypes + a row of ones:
//                        columns 0..K-1 are G P, column K is rowsum(G) — the contraction over prototypes that the
//                        two-kernel version needed a second orientation (and a second evaluation of D, Gl, G) for
//   dZ = 2 (rowsum(G) zc - G Pc) leaves through two writer warps (thread = pixel: coalesced loads of z from L2 and
//   stores of dz).  (A first version kept depth in the lanes: its stores touched 8-32 lines per instruction and the
//   write-out alone took 20-50 k cycles per tile.)
//
// Warp roles (16 warps x 128 registers = the whole register file): 0-7 epilogue (lane quarter = warp % 4, column half =
// warp / 4), 8-11 builder (pixel-side operands of the next tile), 12-13 writer (dZ of the finished tile: pixels in the
// lanes of quarters 0 and 1), 14 spare, 15 MMA issuer.  Every mbarrier completes once per tile or once
// per use of a TMEM buffer, and every waiter takes part in what completes it next (barrier lesson 2).
//
// Shared memory (KP = 64, two prototype blocks, C <= 24): 128-byte-swizzled Zt 16 K + dLt 8 K + G block 0 32 K +
// G block 1 16 K | K-major Pk 48 K (192 rows: block 1 reads rows 128..255, the last 64 alias the next slab and feed
// TMEM lanes nobody reads) + Wt 18 K + Pt 54 K (72 rows: depth, ones, padding) | Zk 16 K + dLk 6 K | own-class
// gradients 8 K + metadata 3 K  = 225 K.

constexpr int FT = 64;   // pixels per tile of the fused backward
constexpr int FU_THREADS = 512;

template <int KP, int NB>
struct FuSmem {
  static constexpr int ROWS = NB == 1 ? 128 : 192;   // prototype rows staged for the K-major A operands
  static constexpr int PR = KP + 8;                  // rows of the transposed prototype operand: depth, ones, padding
  static constexpr int NPROT = NB == 1 ? 128 : 192;  // prototypes the dZ contraction runs over
  static constexpr size_t zt = 0;
  static constexpr size_t dlt = zt + (size_t)(FT / 32) * KP * 128;
  static constexpr size_t gs0 = dlt + (size_t)(FT / 32) * 32 * 128;
  static constexpr size_t gs1 = gs0 + (size_t)4 * FT * 128;
  // two prototype blocks: G of block 1 (64 real prototypes, 2 k-atoms); one block: a second full buffer, the units
  // of consecutive tiles alternate (the dZ^T MMAs of unit g may still read buffer g & 1 while unit g + 1 is written)
  static constexpr size_t GS1_BYTES = (size_t)(NB == 1 ? 4 : 2) * FT * 128;
  static constexpr size_t pk = gs1 + GS1_BYTES;
  static constexpr size_t wt = pk + (size_t)(KP / 4) * ROWS * 16;
  static __host__ __device__ size_t pt(int C8) { return wt + (size_t)(C8 / 4) * ROWS * 16; }
  static __host__ __device__ size_t zk(int C8) { return pt(C8) + (size_t)(NPROT / 4) * PR * 16; }
  static __host__ __device__ size_t dlk(int C8) { return zk(C8) + (size_t)(KP / 4) * FT * 16; }
  static __host__ __device__ size_t ogs(int C8) { return dlk(C8) + (size_t)(C8 / 4) * FT * 16; }
  static __host__ __device__ size_t misc(int C8) { return ogs(C8) + (size_t)2 * FT * JMAX * 4; }
  // misc: x2 [2][FT] f32 | x2 partial [FT] f32 | img [2][FT] i32 | pix [2][FT] i32 | class pixel masks [2][32][2] u32 |
  //       rowsum [FT] f32 | barriers [12] u64 + TMEM base
  static constexpr size_t MISC_BYTES = 512 + 256 + 512 + 512 + 512 + 256 + 128;
  static __host__ __device__ size_t total(int C8) { return misc(C8) + MISC_BYTES; }
};

template <int KP, int NB>
__global__ void __launch_bounds__(FU_THREADS, 1) k_tc_bwd_fused(const TcBwdParams p) {
  using L = FuSmem<KP, NB>;
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr int ROWS = L::ROWS, PR = L::PR, NPROT = L::NPROT, Npad2 = NB * 128, KH = KP / 2;
  const int C8 = p.C8, K = p.K, N = p.N, C = p.C, HW = p.HW;
  float* ZtB = reinterpret_cast<float*>(smem + L::zt);
  float* dLtB = reinterpret_cast<float*>(smem + L::dlt);
  float* Gs0 = reinterpret_cast<float*>(smem + L::gs0);
  float* Gs1 = reinterpret_cast<float*>(smem + L::gs1);
  float* PkA = reinterpret_cast<float*>(smem + L::pk);
  float* WtA = reinterpret_cast<float*>(smem + L::wt);
  float* PtA = reinterpret_cast<float*>(smem + L::pt(C8));
  float* ZkB = reinterpret_cast<float*>(smem + L::zk(C8));
  float* dLkB = reinterpret_cast<float*>(smem + L::dlk(C8));
  float* ogs = reinterpret_cast<float*>(smem + L::ogs(C8));            // [slot][FT px][JMAX]
  uint8_t* misc = smem + L::misc(C8);
  float* x2s = reinterpret_cast<float*>(misc);                          // [slot][FT]
  float* x2p = reinterpret_cast<float*>(misc + 512);                    // [FT] partial |z|^2 of the first depth half
  int* imgs = reinterpret_cast<int*>(misc + 768);                       // [slot][FT]
  int* pixs = reinterpret_cast<int*>(misc + 1280);                      // [slot][FT]
  uint32_t* clsm = reinterpret_cast<uint32_t*>(misc + 1792);            // [slot][32 classes][2 words]
  uint64_t* bars = reinterpret_cast<uint64_t*>(misc + 2560);
  uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(bars + 12);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool kld = p.own_grad != nullptr;
  const bool has_dl = p.dlogits != nullptr;
  const int Mi = (int)p.M;

  if (warp == 15) tmem_alloc(smem_u32(tmem_base_s), 512);
  if (tid == 0) {
    mbar_init(smem_u32(&bars[0]), 128);   // full_k: K-major pixel operands + tile metadata written
    mbar_init(smem_u32(&bars[1]), 1);     // free: every MMA of the tile done (accumulates + dZ^T complete)
    mbar_init(smem_u32(&bars[2]), 1);     // acc[b]: D^T and Gl^T of a unit ready in TMEM buffer b
    mbar_init(smem_u32(&bars[3]), 1);
    mbar_init(smem_u32(&bars[4]), 256);   // gs[b]: G and S of the unit in TMEM buffer b, G in shared memory
    mbar_init(smem_u32(&bars[5]), 256);
    mbar_init(smem_u32(&bars[6]), 128);   // full_t: transposed pixel operands written
    mbar_init(smem_u32(&bars[7]), 1);     // k_free: the tile's last distance MMA has read ZkB / dLkB
    mbar_init(smem_u32(&bars[8]), 64);    // dz_free: dZ of the tile read out of TMEM (two writer warps)
    mbar_init(smem_u32(&bars[9]), 256);   // tile_seen: every epilogue thread has observed full_k of the tile (so the
                                          //   next full_k cannot complete behind a thread's back, and the metadata
                                          //   slot of the tile before is no longer read)
    mbar_init(smem_u32(&bars[10]), 1);    // done: the last MMA of the CTA has retired (completes exactly once)
    mbar_init_fence();
  }
  const int iters = p.num_tiles > (int)blockIdx.x ? (p.num_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const bool is_builder = warp >= 8 && warp < 12;
  const int bt = tid - 256, bpx = bt & (FT - 1), bkh = bt >> 6;   // builder thread = (pixel, depth half)
  float zr[KH];
  // The tile's dlogits block [FT px][C] is contiguous in global memory: every builder thread takes DLV float4s of
  // it (coalesced 16-byte loads) and later scatters them into the two operand layouts — 19 scalar loads per pixel at
  // a 76-byte lane stride kept the load / store unit busy for 5 k cycles per tile.
  constexpr int DLV = (FT * 32 / 4 + 127) / 128;                  // float4s per thread at C = 32
  float4 dlv[DLV];
  uint32_t dle[DLV * 4];                                          // (px << 8) | c of element j of float4 v, or ~0
  if (is_builder) {
#pragma unroll
    for (int v = 0; v < DLV; ++v)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int e = (v * 128 + bt) * 4 + j;
        const int px = e / C, c = e - px * C;
        dle[v * 4 + j] = e < FT * C ? ((uint32_t)px << 8) | (uint32_t)c : 0xffffffffu;
      }
  }
  auto prefetch = [&](int it) {
    const long long m0 = ((long long)blockIdx.x + (long long)it * gridDim.x) * FT;
    const int m = (int)m0 + bpx;
    if (it < iters && m < Mi) {
      const int b = m / HW;
      const float* src = p.z + ((size_t)b * K + (size_t)bkh * KH) * HW + (m - b * HW);
#pragma unroll
      for (int k = 0; k < KH; ++k) zr[k] = (bkh * KH + k < K) ? __ldg(src + (size_t)k * HW) : CENTRE;
    } else {
#pragma unroll
      for (int k = 0; k < KH; ++k) zr[k] = CENTRE;
    }
    const int npx = it < iters ? (int)((p.M - m0) < FT ? (p.M - m0) : FT) : 0;
    const int nfl = npx * C;                                      // floats of this tile's dlogits block
    const float* dl = p.dlogits + (size_t)m0 * C;
    const bool vec = has_dl && (reinterpret_cast<uintptr_t>(dl) & 15u) == 0;
#pragma unroll
    for (int v = 0; v < DLV; ++v) {
      const int e = (v * 128 + bt) * 4;
      float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
      if (has_dl && e < nfl) {
        if (vec && e + 3 < nfl) {
          x = __ldg(reinterpret_cast<const float4*>(dl + e));
        } else {
          x.x = __ldg(dl + e);
          if (e + 1 < nfl) x.y = __ldg(dl + e + 1);
          if (e + 2 < nfl) x.z = __ldg(dl + e + 2);
          if (e + 3 < nfl) x.w = __ldg(dl + e + 3);
        }
      }
      dlv[v] = x;
    }
  };
  if (is_builder) prefetch(0);
  if (warp < 8) {
    // one-time operands (staged by the epilogue warps; the builders have their first tile in flight)
    stage_rows_kmajor<KP, false, 4>(PkA, PkA, p.P, N, K, ROWS, CENTRE, tid, 256);
    stage_wt_kmajor(WtA, p.W, C, N, C8, ROWS, 1.f, has_dl, tid, 256);
    // Pt: rows = depth k (K-major: contraction index = prototype), [NPROT/4][PR][4]; row KP = ones (-> rowsum(G))
    for (int f = tid; f < (NPROT / 4) * PR; f += 256) {
      const int n4 = f / PR, k = f - n4 * PR;
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = n4 * 4 + j;
        float x = 0.f;
        if (n < N) {
          if (k < K) x = rna_tf32_mask(__ldg(p.P + (size_t)n * K + k) - CENTRE);
          else if (k == KP) x = 1.f;
        }
        v[j] = x;
      }
      *reinterpret_cast<float4*>(PtA + (size_t)f * 4) = make_float4(v[0], v[1], v[2], v[3]);
    }
    // the G operands are zeroed once: prototype columns no epilogue warp writes (all-padding warps) must read as 0
    smem_zero4(Gs0, (int)((4 * FT * 128 + L::GS1_BYTES) / 16), tid, 256);          // Gs0 and Gs1 are contiguous
  }
  if (warp >= 12 && warp < 15) {
    smem_zero4(dLkB, (C8 / 4) * FT, tid - 384, 96);                // class columns / rows beyond C stay zero
    smem_zero4(dLtB, (FT / 32) * 32 * 128 / 16, tid - 384, 96);
  }
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  // step path: launched while k_kld may still be running; everything above is independent of it
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;");

  const uint32_t tmem_base = *tmem_base_s;
  // TMEM columns: 0..255 two unit buffers (D^T | Gl^T, 64 + 64 each); dP accumulators [prototype][KP] from 256; dW
  // accumulators [prototype][32] behind them, the two blocks WSTR = 24 columns apart (C8 <= 24: the upper 8 columns of
  // a block are exact zeros, the accumulators start zeroed and every dW MMA accumulates, so a block adds 0.0 to the
  // columns it shares); dZ [pixel][80] at 432..511: depth 0..63 | rowsum(G) | 7 zero columns | 8 unused.  Columns
  // 432..439 are shared by dW block 1's zero columns (which only ever add 0.0) and dZ depth 0..7.
  constexpr int WSTR = NB == 2 ? 24 : 32, ZN = 80;
  const uint32_t tP0 = tmem_base + 256, tW0 = tmem_base + 256 + NB * KP, tZ = tmem_base + 432;
  auto tDb = [&](int b) { return tmem_base + (uint32_t)b * 128u; };
  const uint32_t bar_full = smem_u32(&bars[0]), bar_free = smem_u32(&bars[1]), bar_acc0 = smem_u32(&bars[2]),
                 bar_gs0 = smem_u32(&bars[4]), bar_full_t = smem_u32(&bars[6]), bar_kfree = smem_u32(&bars[7]),
                 bar_dzfree = smem_u32(&bars[8]), bar_seen = smem_u32(&bars[9]), bar_done = smem_u32(&bars[10]);

  if (is_builder) {
    // ================= builder: thread = (pixel, depth half) =================
    for (int it = 0; it < iters; ++it) {
      const long long m0 = ((long long)blockIdx.x + (long long)it * gridDim.x) * FT;
      const int m = (int)m0 + bpx;
      const bool valid = m < Mi;
      const int nfl = (int)((p.M - m0) < FT ? (p.M - m0) : FT) * C;
      const int slot = it & 1;
      uint32_t* clsm_s = clsm + slot * 64;
      float* ogs_s = ogs + slot * FT * JMAX;
      if (warp == 8) DP_PHASE(0, 1 + it * 8);
      // ---- step K: the previous tile's last distance MMA has read ZkB / dLkB (one completion per tile), and every
      //      epilogue thread has left the tile before it (whose metadata slot is written next)
      if (it > 0) {
        mbar_wait(bar_kfree, (uint32_t)(it - 1) & 1u);
        mbar_wait(bar_seen, (uint32_t)(it - 1) & 1u);
      }
      if (warp == 8) DP_PHASE(0, 2 + it * 8);
      if (bt < 64) clsm_s[bt] = 0u;
      if (kld && valid) {
        const float* src = p.own_grad + (size_t)m * JMAX + bkh * (JMAX / 2);
        const uint32_t dst = smem_u32(ogs_s + bpx * JMAX + bkh * (JMAX / 2));
#pragma unroll
        for (int v = 0; v < JMAX / 8; ++v)
          asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst + v * 16), "l"(src + v * 4) : "memory");
      }
      // |z|^2 is one fmaf chain over k = 0 .. K-1, as in the forward kernel (the relu mask must agree with it bit for
      // bit): the first depth half hands its partial sum to the second
      float x2 = 0.f;
      if (bkh == 0) {
#pragma unroll
        for (int k = 0; k < KH; ++k) {
          zr[k] = rna_tf32_mask(zr[k] - CENTRE);
          x2 = fmaf(zr[k], zr[k], x2);
        }
        x2p[bpx] = x2;
      }
      named_bar_sync(2, 128);
      if (bkh == 1) {
        x2 = x2p[bpx];
#pragma unroll
        for (int k = 0; k < KH; ++k) {
          zr[k] = rna_tf32_mask(zr[k] - CENTRE);
          x2 = fmaf(zr[k], zr[k], x2);
        }
        x2s[slot * FT + bpx] = x2;
        const int mylab = (valid && kld) ? p.target[m] : -1;
        if (mylab >= 0 && mylab < 32) atomicOr(&clsm_s[mylab * 2 + (bpx >> 5)], 1u << (bpx & 31));
      }
#pragma unroll
      for (int k4 = 0; k4 < KH / 4; ++k4)
        *reinterpret_cast<float4*>(ZkB + ((size_t)(bkh * (KH / 4) + k4) * FT + bpx) * 4) =
            make_float4(zr[k4 * 4], zr[k4 * 4 + 1], zr[k4 * 4 + 2], zr[k4 * 4 + 3]);
      // dlogits: element e of the tile's [px][C] block -> K-major operand [C8/4][FT][4]; entries beyond the tile's
      // pixels are written as zeros (a shorter last tile must not leave the previous tile's values behind)
#pragma unroll
      for (int v = 0; v < DLV; ++v) {
        float x[4] = {dlv[v].x, dlv[v].y, dlv[v].z, dlv[v].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t pc = dle[v * 4 + j];
          if (pc != 0xffffffffu) {
            const int px = (int)(pc >> 8), c = (int)(pc & 255u);
            x[j] = ((v * 128 + bt) * 4 + j) < nfl ? rna_tf32_fast(x[j]) : 0.f;
            dLkB[((size_t)(c >> 2) * FT + px) * 4 + (c & 3)] = x[j];
          }
        }
      }
      asm volatile("cp.async.wait_all;" ::: "memory");
      fence_async_smem();
      mbar_arrive(bar_full);
      if (warp == 8) DP_PHASE(0, 3 + it * 8);
      // The next tile's global loads go out now: the registers are free (step T below reads this tile's operands back
      // from the K-major copies this thread has just written), and the loads travel while the thread waits.
      prefetch(it + 1);
      if (warp == 8) DP_PHASE(0, 4 + it * 8);
      // ---- step T: the previous tile's accumulate / dZ^T MMAs have retired
      if (it > 0) mbar_wait(bar_free, (uint32_t)(it - 1) & 1u);
      if (warp == 8) DP_PHASE(0, 5 + it * 8);
#pragma unroll
      for (int k4 = 0; k4 < KH / 4; ++k4) {
        const float4 v = *reinterpret_cast<const float4*>(ZkB + ((size_t)(bkh * (KH / 4) + k4) * FT + bpx) * 4);
        const int k = bkh * KH + k4 * 4;
        ZtB[off_sw128(k, bpx, KP)] = v.x;
        ZtB[off_sw128(k + 1, bpx, KP)] = v.y;
        ZtB[off_sw128(k + 2, bpx, KP)] = v.z;
        ZtB[off_sw128(k + 3, bpx, KP)] = v.w;
      }
#pragma unroll
      for (int v = 0; v < DLV; ++v) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t pc = dle[v * 4 + j];
          if (pc != 0xffffffffu) {
            const int px = (int)(pc >> 8), c = (int)(pc & 255u);
            dLtB[off_sw128(c, px, 32)] = dLkB[((size_t)(c >> 2) * FT + px) * 4 + (c & 3)];
          }
        }
      }
      fence_async_smem();
      mbar_arrive(bar_full_t);
      if (warp == 8) DP_PHASE(0, 6 + it * 8);
    }
  } else if (warp >= 12 && warp < 14) {
    // ================= writer: dZ = 2 (rowsum(G) zc - G Pc) of the finished tile, thread = pixel =================
    // dZ sits in TMEM with pixels in lanes 0..63 (M = 128 MMA whose upper 64 rows are over-read garbage): lane quarters
    // 0 and 1.  Loads of z and stores of dz are coalesced along the pixels, as in the forward kernel.
    const int q = warp - 12, px = q * 32 + lane;
    const uint32_t taddr = tZ + ((uint32_t)(q * 32) << 16);
    for (int it = 0; it < iters; ++it) {
      const long long m0 = ((long long)blockIdx.x + (long long)it * gridDim.x) * FT;
      const int m = (int)m0 + px;
      const bool valid = m < Mi && !(p.exp_flags & 1);
      const int b = valid ? m / HW : 0;
      const size_t base = ((size_t)b * K) * HW + (valid ? m - b * HW : 0);
      mbar_wait(bar_free, (uint32_t)it & 1u);
      fence_after();
      if (warp == 12) DP_PHASE(0, 7 + it * 8);
      // the accumulator is copied to registers and handed back at once (the next tile's dZ MMAs wait for it); the
      // arithmetic, the loads of z and the stores follow off the critical path
      uint32_t r[KP / 16][16], rsr[16];
#pragma unroll
      for (int ch = 0; ch < KP / 16; ++ch) tmem_ld16(taddr + ch * 16, r[ch]);
      tmem_ld16(taddr + KP, rsr);
      tmem_ld_wait();
      fence_before();
      mbar_arrive(bar_dzfree);
      const float rs = __uint_as_float(rsr[0]);
#pragma unroll
      for (int ch = 0; ch < KP / 16; ++ch) {
        float zv[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int k = ch * 16 + i;
          zv[i] = (valid && k < K) ? __ldg(p.z + base + (size_t)k * HW) : CENTRE;
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int k = ch * 16 + i;
          if (valid && k < K) {
            float gz = 2.f * fmaf(rs, zv[i] - CENTRE, -__uint_as_float(r[ch][i]));
            if (p.dz_sigmoid) gz *= zv[i] * (1.f - zv[i]);
            p.dz[base + (size_t)k * HW] = gz;
          }
        }
      }
    }
  } else if (warp == 14) {
    // spare warp (no role since the dZ accumulator has pixels in its lanes: two writer warps cover them)
  } else if (warp == 15) {
    // ================= MMA issuer =================
    // all 32 lanes run the control flow (warp-uniform); one elected lane issues each MMA / commit (tc_common.cuh)
    if (iters > 0) {
      const uint32_t idesc_t = make_idesc_tf32(128, FT), idesc_p = make_idesc_tf32(128, KP),
                     idesc_w = make_idesc_tf32(128, 32), idesc_z = make_idesc_tf32(128, ZN);
      const int ksteps = (K + 7) >> 3, csteps = C8 >> 3;
      const int total_units = iters * NB;
      // every descriptor is a base plus a compile-time multiple of 16 bytes in the start-address field: the issuing
      // thread spent ~80 cycles per MMA building descriptors before
      const uint64_t d_zt = make_desc_sw128(smem_u32(ZtB)), d_dlt = make_desc_sw128(smem_u32(dLtB));
      const uint64_t d_pt = make_desc(smem_u32(PtA), PR * 16, 128);
      const uint64_t d_g0 = make_desc_sw128(smem_u32(Gs0)), d_g1 = make_desc_sw128(smem_u32(Gs1));
      // unit g = it * NB + blk uses TMEM buffer g & 1
      auto issue_d = [&](int g) {
        const int blk = g % NB, b = g & 1;
        const uint32_t td = tDb(b);
        // (descriptors built from the byte addresses here: the same operands addressed by adding to precomputed 64-bit
        // descriptors, as the accumulate MMAs below do, gave wrong distance / Gl tiles from the second tile of a CTA
        // on — bisected on the GPU, cause not understood; this form is the one the two-kernel backward uses)
        const uint32_t a0b = smem_u32(PkA) + blk * 128 * 16, b0b = smem_u32(ZkB);
        const uint32_t w0b = smem_u32(WtA) + blk * 128 * 16, l0b = smem_u32(dLkB);
        // One accumulation chain after the other: switching the accumulator between consecutive MMAs measured SLOWER
        // (interleaved D / Gl chains: 96 -> 114 us at B = 16) — the tensor pipe streams a chain into one accumulator.
        for (int ks = 0; ks < ksteps; ++ks)
          mma_ss_w(td, make_desc(a0b + ks * 2 * ROWS * 16, ROWS * 16, 128), make_desc(b0b + ks * 2 * FT * 16, FT * 16, 128),
                   idesc_t, ks > 0 ? 1u : 0u);
        for (int ks = 0; ks < csteps; ++ks)
          mma_ss_w(td + FT, make_desc(w0b + ks * 2 * ROWS * 16, ROWS * 16, 128), make_desc(l0b + ks * 2 * FT * 16, FT * 16, 128),
                   idesc_t, ks > 0 ? 1u : 0u);
        mma_commit_w(bar_acc0 + b * 8);
        if (blk == NB - 1) mma_commit_w(bar_kfree);
      };
      uint32_t gcnt[2] = {0u, 0u};
      int fullk_seen = 0;                      // tiles whose full_k completion has been observed
      auto need_tile = [&](int it) {
        while (fullk_seen <= it) {
          mbar_wait(bar_full, (uint32_t)fullk_seen & 1u);
          ++fullk_seen;
        }
        fence_after();
      };
      need_tile(0);
      issue_d(0);
      if (total_units > 1) {
        need_tile(1 / NB);
        issue_d(1);
      }
      for (int g = 0; g < total_units; ++g) {
        const int it = g / NB, blk = g % NB, b = g & 1;
        mbar_wait(bar_gs0 + b * 8, gcnt[b] & 1u);
        ++gcnt[b];
        if (blk == 0) {
          mbar_wait(bar_full_t, (uint32_t)it & 1u);
          if (it > 0) mbar_wait(bar_dzfree, (uint32_t)(it - 1) & 1u);
        }
        fence_after();
        DP_PHASE(1, 1 + it * 8 + blk * 4);
        const uint32_t td = tDb(b), tp = tP0 + blk * KP, tw = tW0 + blk * WSTR;
        const uint32_t acc0 = it > 0 ? 1u : 0u;
        // three accumulation chains (dP, dW, dZ), one after the other
        const uint64_t pa = d_pt + (uint64_t)(blk * 32 * PR);
        const uint64_t gb = (NB == 1 ? (g & 1) : blk) == 0 ? d_g0 : d_g1;
        const uint32_t accz = blk > 0 ? 1u : 0u;
        const bool with_z = !(p.exp_flags & 2);
        constexpr int NK0 = (NPROT < 128 ? NPROT : 128) / 8;       // dZ k-steps of block 0 (block 1: the rest)
        const int nkz = blk == 0 ? NK0 : (NPROT - 128) / 8;
        auto z_step = [&](int ks) {
          if (with_z && ks < nkz)
            mma_ss_w(tZ, gb + (uint64_t)(((ks >> 2) * FT * 128 + (ks & 3) * 32) >> 4), pa + (uint64_t)(ks * 2 * PR), idesc_z,
                     ks > 0 ? 1u : accz);
        };
#pragma unroll
        for (int ks = 0; ks < FT / 8; ++ks)
          mma_ts_w(tp, td + ks * 8, d_zt + (uint64_t)(((ks >> 2) * KP * 128 + (ks & 3) * 32) >> 4), idesc_p, ks > 0 ? 1u : acc0);
#pragma unroll
        for (int ks = 0; ks < FT / 8; ++ks)
          mma_ts_w(tw, td + FT + ks * 8, d_dlt + (uint64_t)(((ks >> 2) * 32 * 128 + (ks & 3) * 32) >> 4), idesc_w, 1u);
#pragma unroll
        for (int ks = 0; ks < NK0; ++ks) z_step(ks);
        if (blk == NB - 1) mma_commit_w(bar_free);
        if (g == total_units - 1) mma_commit_w(bar_done);
        DP_PHASE(1, 2 + it * 8 + blk * 4);
        if (g + 2 < total_units) {
          need_tile((g + 2) / NB);
          issue_d(g + 2);
        }
        DP_PHASE(1, 3 + it * 8 + blk * 4);
      }
    }
  } else {
    // ================= epilogue: thread = (prototype lane, pixel-column half) =================
    const int half = warp >> 2, q = warp & 3, lane_n = (q << 5) + lane;
    const uint32_t lane_field = (uint32_t)(q * 32) << 16;
    float p2[NB], cs[NB];
    int pcl[NB], prk[NB];
#pragma unroll
    for (int blk = 0; blk < NB; ++blk) {
      const int n = blk * 128 + lane_n;
      float a = 0.f;
      if (n < ROWS) {
        for (int k4 = 0; k4 < KP / 4; ++k4) {
          const float4 v = *reinterpret_cast<const float4*>(PkA + ((size_t)k4 * ROWS + n) * 4);
          a = fmaf(v.x, v.x, a); a = fmaf(v.y, v.y, a); a = fmaf(v.z, v.z, a); a = fmaf(v.w, v.w, a);
        }
      }
      p2[blk] = a;
      cs[blk] = 0.f;
      pcl[blk] = -1;
      prk[blk] = 0;
      if (kld && n < N) { pcl[blk] = p.proto_class[n]; prk[blk] = p.proto_rank[n]; }
      if (pcl[blk] < 0 || pcl[blk] >= 32 || prk[blk] < 0) { pcl[blk] = -1; prk[blk] = 0; }
    }
    // byte offset of this lane's element in row (c & 7) of the swizzled G operand (rows = pixels, k = prototype)
    uint32_t gofs[8];
#pragma unroll
    for (int j = 0; j < 8; ++j)
      gofs[j] = (uint32_t)q * (FT * 128) + (uint32_t)j * 128 + ((((uint32_t)lane >> 2) ^ (uint32_t)j) << 4) + ((uint32_t)lane & 3u) * 4u;
    const float eps = p.eps, epsm1 = p.eps - 1.f;
    const bool act_log = p.act_mode == ADLM_ACT_LOG;
    {
      // the dW accumulators start at zero (their MMAs always accumulate, see the TMEM map above); ordered before the
      // first dW MMA by this thread's first arrival on gs[]
      uint32_t zz[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) zz[i] = 0u;
      for (int c = half * NB * 16; c < (half + 1) * NB * 16; c += 16) tmem_st16(tW0 + lane_field + c, zz);
      tmem_st_wait();
    }
    uint32_t acnt[2] = {0u, 0u};
    for (int it = 0; it < iters; ++it) {
      mbar_wait(bar_full, (uint32_t)it & 1u);
      mbar_arrive(bar_seen);
      const int slot = it & 1;
      const float* x2s_s = x2s + slot * FT;
      const uint32_t* clsm_s = clsm + slot * 64;
      const float* ogs_s = ogs + slot * FT * JMAX;
#pragma unroll
      for (int blk = 0; blk < NB; ++blk) {
        const int b = (it * NB + blk) & 1;
        const uint32_t tD1 = tDb(b), tGl = tDb(b) + FT;
        if (warp == 0) DP_PHASE(2, 1 + it * 8 + blk * 4);
        mbar_wait(bar_acc0 + b * 8, acnt[b] & 1u);
        if (warp == 0) DP_PHASE(2, 2 + it * 8 + blk * 4);
        ++acnt[b];
        fence_after();
        if (blk * 128 + q * 32 < N) {       // warps whose 32 prototypes are all padding skip the math
          const uint32_t gs_base = smem_u32((NB == 1 ? (it & 1) : blk) == 0 ? Gs0 : Gs1);
#pragma unroll 1
          for (int ch = 0; ch < 2; ++ch) {
            const int c0 = half * 32 + ch * 16;       // pixel column of the tile
            uint32_t ra[16], rg[16], og[16], os[16];
            tmem_ld16(tD1 + lane_field + c0, ra);
            tmem_ld16(tGl + lane_field + c0, rg);
            float xp[16];
#pragma unroll
            for (int v4 = 0; v4 < 4; ++v4) {
              const float4 qv = *reinterpret_cast<const float4*>(x2s_s + c0 + v4 * 4);
              xp[v4 * 4] = qv.x + p2[blk]; xp[v4 * 4 + 1] = qv.y + p2[blk]; xp[v4 * 4 + 2] = qv.z + p2[blk]; xp[v4 * 4 + 3] = qv.w + p2[blk];
            }
            const uint32_t hit = pcl[blk] >= 0 ? ((clsm_s[pcl[blk] * 2 + (c0 >> 5)] >> (c0 & 31)) & 0xffffu) : 0u;
            tmem_ld_wait();
            float gvv[16];
            uint32_t pos = 0u;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float pre = fmaf(-2.f, __uint_as_float(ra[i]), xp[i]);
              const float d = fmaxf(pre, 0.f);
              float gv = __uint_as_float(rg[i]);
              float sv;
              if (act_log) {
                const float u = d + 1.f, v = d + eps;
                sv = lg2_approx(u) - lg2_approx(v);             // log2 units; ln 2 is folded into the flush
                gv *= epsm1 * rcp_approx(u * v);
              } else {
                sv = -d;
                gv = -gv;
              }
              pos |= (pre > 0.f ? 1u : 0u) << i;
              gvv[i] = gv;
              os[i] = __float_as_uint(rna_tf32_fast(sv));
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
            }
            tmem_st16(tD1 + lane_field + c0, og);
            tmem_st16(tGl + lane_field + c0, os);
            if ((NB == 1 || blk == 0 || q < 2) && !(p.exp_flags & 4)) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const int c = c0 + i;
                asm volatile("st.shared.b32 [%0], %1;" ::"r"(gs_base + gofs[c & 7] + (uint32_t)(c >> 3) * 1024u), "r"(og[i]) : "memory");
              }
            }
          }
          tmem_st_wait();
        }
        fence_before();
        fence_async_smem();
        mbar_arrive(bar_gs0 + b * 8);
        if (warp == 0) DP_PHASE(2, 3 + it * 8 + blk * 4);
      }
    }
    // ---- flush this CTA's partial sums ----
    if (iters > 0) {
      mbar_wait(bar_done, 0u);
      fence_after();
    }
    float* css = ZkB;                                             // every MMA has retired: operand buffers are scratch
#pragma unroll
    for (int blk = 0; blk < NB; ++blk) css[(half * 2 + blk) * 128 + lane_n] = cs[blk];
    if (NB == 1) css[(half * 2 + 1) * 128 + lane_n] = 0.f;
    named_bar_sync(1, 256);
    const float s_scale = act_log ? LN2 : 1.f;
    float* part = p.partials + (size_t)blockIdx.x * ((size_t)Npad2 * KP + (size_t)Npad2 * 32);
    constexpr int HK = KP / 2, PITCH = HK + 4;
    static_assert(8 * 32 * PITCH * 4 <= 4 * FT * 128 + L::GS1_BYTES, "transpose scratch exceeds the G operand buffers");
    float* scr = Gs0 + (size_t)warp * 32 * PITCH;
    for (int blk = 0; blk < NB; ++blk) {
      const int n = blk * 128 + lane_n, n_base = blk * 128 + q * 32;
      const float cst = css[blk * 128 + lane_n] + css[(2 + blk) * 128 + lane_n];
      if (n_base < N) {
#pragma unroll
        for (int ch = 0; ch < KP / 32; ++ch) {
          const int k0 = half * HK + ch * 16;
          uint32_t r[16];
          if (iters > 0) {
            tmem_ld16(tP0 + blk * KP + lane_field + k0, r);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) r[i] = 0u;
          }
#pragma unroll
          for (int i4 = 0; i4 < 4; ++i4) {
            float o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int k = k0 + i4 * 4 + j;
              const float pc = n < ROWS ? PkA[(size_t)(k >> 2) * ROWS * 4 + n * 4 + (k & 3)] : 0.f;
              o[j] = 2.f * (cst * pc - __uint_as_float(r[i4 * 4 + j]));
            }
            *reinterpret_cast<float4*>(scr + lane * PITCH + ch * 16 + i4 * 4) = make_float4(o[0], o[1], o[2], o[3]);
          }
        }
        __syncwarp();
        {
          constexpr int V = HK / 4, RPI = 32 / V;
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
        if (iters > 0) {
          tmem_ld16(tW0 + blk * WSTR + lane_field + half * 16, r);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) r[i] = 0u;
        }
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
  __syncthreads();
  if (warp == 15) tmem_dealloc(tmem_base, 512);
}
