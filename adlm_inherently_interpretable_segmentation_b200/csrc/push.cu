// Push: nearest-patch search of segmentation/push.py as streaming kernels.
//
//   k_patch_class_mask  push.py:201-230  full-resolution labels -> per-patch class bitmask
//   k_simt_push_scan    push.py:249-280  class-filtered per-prototype argmin, fp32 arithmetic
//   k_push_gather       push.py:268-280, :142-145  winners' feature vectors
//
// The reference copies the (h,w,N) distance map and the features of every image to the host and
// searches them in Python; here the distance never leaves the SM: each CTA keeps a running
// (distance bits << 32 | global patch id) minimum per prototype in shared memory and folds it into
// the global array with one 64-bit atomicMin per prototype.
#include <cstdlib>
#include "common.cuh"

namespace adlm {

namespace {

// One CTA per (patch row, image): OR the class bits of every label pixel whose row maps to this patch row into
// a shared-memory row of w masks.  Label rows are streamed 16 pixels (one 16-byte load) per thread; the 16
// pixels fall into at most a few patches, so their bits are combined in registers and flushed with one
// shared-memory atomicOr per patch (skipped when the bits are already there).
template <bool VEC>
__global__ void __launch_bounds__(256)
k_patch_class_mask(const uint8_t* __restrict__ labels, int H, int Wimg, const int32_t* __restrict__ row_lut,
                   const int32_t* __restrict__ col_lut, int h, int w, uint32_t* __restrict__ mask_out) {
  extern __shared__ uint32_t smask[];
  __shared__ int s_rows[2];
  const int pi = blockIdx.x, img = blockIdx.y;
  for (int j = threadIdx.x; j < w; j += blockDim.x) smask[j] = 0u;
  if (threadIdx.x < 2) s_rows[threadIdx.x] = 0;
  __syncthreads();
  const uint8_t* lab = labels + (size_t)img * H * Wimg;
  // first / one-past-last label row of this patch row: the LUT is monotone, so the CTA looks at every row once, in
  // parallel (one round of cached loads; the two binary searches this replaces were 22 dependent round trips in a
  // CTA whose whole payload is one batch of label loads)
  for (int r = threadIdx.x; r < H; r += blockDim.x) {
    if (__ldg(row_lut + r) == pi) {
      if (r == 0 || __ldg(row_lut + r - 1) != pi) s_rows[0] = r;
      if (r == H - 1 || __ldg(row_lut + r + 1) != pi) s_rows[1] = r + 1;
    }
  }
  __syncthreads();
  const int r0 = s_rows[0], r1 = s_rows[1];
  if (VEC) {
    const int vpr = Wimg >> 4;                      // 16-pixel vectors per row
    const int nvec = (r1 - r0) * vpr;
    const uint4* base = reinterpret_cast<const uint4*>(lab + (size_t)r0 * Wimg);
    auto consume = [&](int vec, const uint4 v) {
      const int cv = vec % vpr;
      // label maps are piecewise constant: 16 equal labels need one bit and the patch range of the run
      // (the column LUT is monotone with steps <= 1 when patches are at least one pixel wide, w <= Wimg)
      if (w <= Wimg && v.y == v.x && v.z == v.x && v.w == v.x && v.x == (v.x & 0xffu) * 0x01010101u) {
        const unsigned l = v.x & 0xffu;
        if (l - 1u < 32u) {
          const uint32_t bit = 1u << (l - 1u);
          const int pj1 = __ldg(col_lut + cv * 16 + 15);
          for (int pj = __ldg(col_lut + cv * 16); pj <= pj1; ++pj)
            if ((smask[pj] & bit) == 0u) atomicOr(&smask[pj], bit);
        }
        return;
      }
      const int4* lut4 = reinterpret_cast<const int4*>(col_lut + cv * 16);
      const uint32_t words[4] = {v.x, v.y, v.z, v.w};
      int cur = -1;
      uint32_t bits = 0u;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int4 pj4 = __ldg(lut4 + q);
        const int pjs[4] = {pj4.x, pj4.y, pj4.z, pj4.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const unsigned l = (words[q] >> (8 * e)) & 0xffu;
          if (pjs[e] != cur) {
            if (bits != 0u && (smask[cur] & bits) != bits) atomicOr(&smask[cur], bits);
            cur = pjs[e];
            bits = 0u;
          }
          if (l - 1u < 32u) bits |= 1u << (l - 1u);
        }
      }
      if (bits != 0u && (smask[cur] & bits) != bits) atomicOr(&smask[cur], bits);
    };
    for (int v0 = threadIdx.x; v0 < nvec; v0 += 4 * 256) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int vec = v0 + u * 256;
        v[u] = vec < nvec ? __ldg(base + vec) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int vec = v0 + u * 256;
        if (vec < nvec) consume(vec, v[u]);
      }
    }
  } else {
    const int n = (r1 - r0) * Wimg;
    for (int idx = threadIdx.x; idx < n; idx += blockDim.x) {
      const int r = r0 + idx / Wimg, j = idx - (idx / Wimg) * Wimg;
      const unsigned l = lab[(size_t)r * Wimg + j];
      if (l > 0u && l <= 32u) {
        const uint32_t bit = 1u << (l - 1u);
        const int pj = col_lut[j];
        if ((smask[pj] & bit) == 0u) atomicOr(&smask[pj], bit);
      }
    }
  }
  __syncthreads();
  uint32_t* out = mask_out + ((size_t)img * h + pi) * w;
  for (int j = threadIdx.x; j < w; j += blockDim.x) out[j] = smask[j];
}

// The same, one THREAD per patch: no shared memory, no atomics, no local arrays.  A patch of the reference's maps is
// ~8 x 8 label pixels (1024 x 2048 labels -> 129 x 257 patches); a thread reads, for each of its label rows, the one
// or two aligned 8-byte words that cover its columns and ORs the class bits of the bytes inside its range (one
// compare when all of them are equal, which piecewise-constant label maps make the common case).  The 32 patches
// of a warp cover ~256 contiguous bytes per row, so the loads coalesce; neighbours' overlapping words hit L1.
// Row / column ranges come from the monotone LUTs by a local search around the proportional guess; a CTA covers
// MASK_PR patch rows so a thread finds its column range once.  Needs Wimg % 8 == 0 and 8-byte aligned labels.
// Measured (tools/mask_bench.py, 512 label maps of 1024 x 2048): 446 us = 2.4 TB/s whatever the label granularity;
// the one-CTA-per-patch-row kernel above (ADLM_MASK_CTA_KERNEL=1) does 405 us on 32-pixel label blocks but 880 us on
// 8-pixel blocks (its slow path and shared-memory atomics).  Both are instruction-bound, not HBM-bound: a variant with
// all 16 loads of a patch in one batch and 32-bit arithmetic needed 83 registers and ran at 787 us.
__device__ __forceinline__ int lut_lower_bound(const int32_t* __restrict__ lut, int n, int target, int guess) {
  // first index i with lut[i] >= target (lut non-decreasing); `guess` only sets where the search starts
  int g = guess < 0 ? 0 : (guess > n ? n : guess);
  while (g > 0 && __ldg(lut + g - 1) >= target) --g;
  while (g < n && __ldg(lut + g) < target) ++g;
  return g;
}

constexpr int MASK_PR = 8;          // patch rows per CTA: a thread's column range is found once and used for all of them

__global__ void __launch_bounds__(320)
k_patch_class_mask_thread(const uint8_t* __restrict__ labels, int H, int Wimg, const int32_t* __restrict__ row_lut,
                          const int32_t* __restrict__ col_lut, int h, int w, uint32_t* __restrict__ mask_out) {
  const int pi0 = blockIdx.x * MASK_PR, img = blockIdx.y;
  const int pi1 = pi0 + MASK_PR < h ? pi0 + MASK_PR : h;
  const uint8_t* lab = labels + (size_t)img * H * Wimg;
  const int rstep = H / h;
  for (int pj = threadIdx.x; pj < w; pj += blockDim.x) {
    const int c0 = lut_lower_bound(col_lut, Wimg, pj, (int)(((unsigned)pj * (unsigned)Wimg) / (unsigned)w));
    const int c1 = lut_lower_bound(col_lut, Wimg, pj + 1, c0 + Wimg / w);
    const int a0 = c0 & ~7;
    int r0 = lut_lower_bound(row_lut, H, pi0, (int)(((unsigned)pi0 * (unsigned)H) / (unsigned)h));
    for (int pi = pi0; pi < pi1; ++pi) {
      const int r1 = lut_lower_bound(row_lut, H, pi + 1, r0 + rstep);
      uint32_t bits = 0u;
      for (int a = a0; a < c1; a += 8) {                        // aligned words that touch [c0, c1)
        const int lo = c0 > a ? c0 - a : 0, hi = c1 - a < 8 ? c1 - a : 8;
        const unsigned long long M = (hi - lo == 8) ? ~0ull : (((1ull << ((hi - lo) * 8)) - 1ull) << (lo * 8));
        for (int rb = r0; rb < r1; rb += 8) {                   // eight rows in flight
          unsigned long long wv[8];
#pragma unroll
          for (int u = 0; u < 8; ++u)
            wv[u] = rb + u < r1 ? __ldg(reinterpret_cast<const unsigned long long*>(lab + (size_t)(rb + u) * Wimg + a)) : 0ull;
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            if (rb + u >= r1) break;
            const unsigned l0 = (unsigned)(wv[u] >> (lo * 8)) & 0xffu;
            if (((wv[u] ^ (l0 * 0x0101010101010101ull)) & M) == 0ull) {
              if (l0 - 1u < 32u) bits |= 1u << (l0 - 1u);
            } else {
              for (int bq = lo; bq < hi; ++bq) {
                const unsigned l = (unsigned)(wv[u] >> (bq * 8)) & 0xffu;
                if (l - 1u < 32u) bits |= 1u << (l - 1u);
              }
            }
          }
        }
      }
      mask_out[((size_t)img * h + pi) * w + pj] = bits;
      r0 = r1;
    }
  }
}

constexpr int TM = 64, TN = 64, NT = 256, PS = TN + 4;

// Same tiling as simt_head.cu's forward; the epilogue is the masked running minimum.
__global__ void __launch_bounds__(NT)
k_simt_push_scan(const float* __restrict__ z, int K, int HW, const uint32_t* __restrict__ mask,
                 const float* __restrict__ P, int N, const int32_t* __restrict__ proto_class,
                 unsigned long long patch_base, unsigned long long* __restrict__ packed_min,
                 int tiles_per_img, int per_image) {
  extern __shared__ __align__(16) float smem[];
  const int Kq = (K + 63) & ~63;
  float* zs = smem;
  float* ps = zs + Kq * TM;
  float* x2s = ps + Kq * PS;
  float* p2s = x2s + TM;
  unsigned long long* best = reinterpret_cast<unsigned long long*>(p2s + TN);   // [TN]
  uint32_t* ms = reinterpret_cast<uint32_t*>(best + TN);                        // [TM]
  int* pcs = reinterpret_cast<int*>(ms + TM);                                   // [TN]

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int img = blockIdx.x / tiles_per_img, p0 = (blockIdx.x % tiles_per_img) * TM;
  const float* zb = z + (size_t)img * K * HW;
  for (int idx = tid; idx < Kq * TM; idx += NT) {
    const int k = idx >> 6, px = idx & 63, p = p0 + px;
    zs[idx] = (k < K && p < HW) ? __ldg(zb + (size_t)k * HW + p) : 0.f;
  }
  if (tid < TM) ms[tid] = (p0 + tid < HW) ? (mask ? mask[(size_t)img * HW + p0 + tid] : 0xffffffffu) : 0u;
  __syncthreads();
  if (tid < TM) {
    float a = 0.f;
    for (int k = 0; k < K; ++k) { const float v = zs[k * TM + tid]; a = fmaf(v, v, a); }
    x2s[tid] = a;
  }
  uint32_t any = 0u;
  for (int i = 0; i < TM; ++i) any |= ms[i];
  if (any == 0u) return;   // no labelled pixel in this tile (uniform across the CTA)

  for (int n0 = 0; n0 < N; n0 += TN) {
    __syncthreads();
    const int kblocks = Kq >> 3;
    for (int idx = tid; idx < TN * Kq; idx += NT) {
      const int kk = idx & 7, nn = (idx >> 3) & 3, rest = idx >> 5;
      const int kb = rest % kblocks, nb = rest / kblocks;
      const int n = nb * 4 + nn, k = kb * 8 + kk;
      ps[k * PS + n] = (n0 + n < N && k < K) ? __ldg(P + (size_t)(n0 + n) * K + k) : 0.f;
    }
    if (tid < TN) {
      best[tid] = ADLM_PUSH_NONE;
      pcs[tid] = (n0 + tid < N) ? (proto_class ? proto_class[n0 + tid] : 0) : -1;
    }
    __syncthreads();
    if (tid >= 64 && tid < 128) {
      const int n = tid - 64;
      float a = 0.f;
      for (int k = 0; k < K; ++k) { const float v = ps[k * PS + n]; a = fmaf(v, v, a); }
      p2s[n] = a;
    }
    float acc[4][4];
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
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int nl = ty * 4 + j;
      const int pc = pcs[nl];
      unsigned long long key = ADLM_PUSH_NONE;
      if (pc >= 0 && pc < 32) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int px = tx + 16 * i;
          if ((ms[px] >> pc) & 1u) {
            const float d = fmaxf(x2s[px] + fmaf(-2.f, acc[i][j], p2s[nl]), 0.f);
            const unsigned long long kk =
                ((unsigned long long)__float_as_uint(d) << 32) |
                (per_image ? (unsigned long long)(p0 + px)
                           : patch_base + (unsigned long long)img * HW + (unsigned long long)(p0 + px));
            key = kk < key ? kk : key;
          }
        }
      }
      // the 16 lanes sharing ty hold the same prototype: shuffle-min across them
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
        key = other < key ? other : key;
      }
      if (tx == 0 && key != ADLM_PUSH_NONE) best[nl] = key;   // one writer per prototype
    }
    __syncthreads();
    if (tid < TN && n0 + tid < N && best[tid] != ADLM_PUSH_NONE)
      atomicMin(packed_min + (per_image ? (size_t)img * N : (size_t)0) + n0 + tid, best[tid]);
  }
}

__global__ void k_push_gather(const float* __restrict__ z, int n_img, int K, int HW,
                              unsigned long long patch_base, const unsigned long long* __restrict__ packed,
                              int N, float* __restrict__ out) {
  const int n = blockIdx.x;
  const unsigned long long key = packed[n];
  if (key == ADLM_PUSH_NONE) return;
  const unsigned long long id = key & 0xffffffffULL;
  if (id < patch_base || id >= patch_base + (unsigned long long)n_img * HW) return;
  const unsigned long long local = id - patch_base;
  const size_t img = local / HW, p = local % HW;
  for (int k = threadIdx.x; k < K; k += blockDim.x) out[(size_t)n * K + k] = z[(img * K + k) * HW + p];
}

// Legacy min-pool classification branch (model.py:284-299): the (B,N) minima come from the per-image scan;
// these two kernels are its (B*N*K)-sized ends.
// forward tail: unpack the per-image minima into min_distances [B,N] (+inf -> never happens without a mask)
__global__ void k_min_pool_unpack(const unsigned long long* __restrict__ packed, int total,
                                  float* __restrict__ min_dist, int32_t* __restrict__ argmin) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const unsigned long long key = packed[i];
  const bool none = key == ADLM_PUSH_NONE;
  min_dist[i] = none ? __int_as_float(0x7f800000) : __uint_as_float((uint32_t)(key >> 32));
  if (argmin) argmin[i] = none ? -1 : (int32_t)(key & 0xffffffffULL);
}

// backward: the gradient of min_distances[b,n] lands on the single patch that attains the minimum
// (max_pool2d backward) and passes the relu only where d > 0 (model.py:218):
//   dz[b,:,patch] += 2 g (z - p),  dP[n,:] -= 2 g (z - p)
__global__ void k_min_pool_bwd(const float* __restrict__ z, int K, int HW, const float* __restrict__ P, int N,
                               const unsigned long long* __restrict__ packed, const float* __restrict__ gmin,
                               float* __restrict__ dz, float* __restrict__ dP) {
  const int n = blockIdx.x, b = blockIdx.y;
  const unsigned long long key = packed[(size_t)b * N + n];
  if (key == ADLM_PUSH_NONE) return;
  const float d = __uint_as_float((uint32_t)(key >> 32));
  const float g = gmin[(size_t)b * N + n];
  if (!(d > 0.f) || g == 0.f) return;
  const size_t patch = (size_t)(key & 0xffffffffULL);
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    const size_t zi = ((size_t)b * K + k) * HW + patch;
    const float t = 2.f * g * (__ldg(z + zi) - __ldg(P + (size_t)n * K + k));
    if (dz) atomicAdd(dz + zi, t);
    if (dP) atomicAdd(dP + (size_t)n * K + k, -t);
  }
}

}  // namespace

int min_pool_unpack_impl(const unsigned long long* packed, int B, int N, float* min_dist, int32_t* argmin,
                         cudaStream_t st) {
  const int total = B * N;
  k_min_pool_unpack<<<cdiv(total, 256), 256, 0, st>>>(packed, total, min_dist, argmin);
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

int min_pool_bwd_impl(const float* z, int B, int K, int HW, const float* P, int N,
                      const unsigned long long* packed, const float* gmin, float* dz, float* dP, cudaStream_t st) {
  dim3 grid(N, B);
  k_min_pool_bwd<<<grid, 64, 0, st>>>(z, K, HW, P, N, packed, gmin, dz, dP);
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

int patch_class_mask_impl(const uint8_t* labels, int n_img, int H, int Wimg, const int32_t* row_lut,
                          const int32_t* col_lut, int h, int w, uint32_t* mask_out, cudaStream_t st) {
  dim3 grid(h, n_img);
  KernelTimer timer(ADLM_K_PATCH_MASK, st);
  // 16-byte label loads need Wimg % 16 == 0 and 16-byte aligned rows / LUT
  const bool vec = (Wimg % 16 == 0) && ((reinterpret_cast<uintptr_t>(labels) & 15u) == 0) &&
                   ((reinterpret_cast<uintptr_t>(col_lut) & 15u) == 0);
  const bool per_thread = (Wimg % 8 == 0) && ((reinterpret_cast<uintptr_t>(labels) & 7u) == 0) && h <= H && w <= Wimg &&
                          getenv("ADLM_MASK_CTA_KERNEL") == nullptr;
  if (per_thread) {
    const int threads = w >= 320 ? 320 : ((w + 31) / 32) * 32;
    grid.x = (h + MASK_PR - 1) / MASK_PR;
    k_patch_class_mask_thread<<<grid, threads, 0, st>>>(labels, H, Wimg, row_lut, col_lut, h, w, mask_out);
  } else if (vec) k_patch_class_mask<true><<<grid, 256, sizeof(uint32_t) * w, st>>>(labels, H, Wimg, row_lut, col_lut, h, w, mask_out);
  else     k_patch_class_mask<false><<<grid, 256, sizeof(uint32_t) * w, st>>>(labels, H, Wimg, row_lut, col_lut, h, w, mask_out);
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

int simt_push_scan(const float* z, int n_img, int K, int HW, const uint32_t* mask, const float* P, int N,
                   const int32_t* proto_class, long long patch_base, unsigned long long* packed_min,
                   int per_image, cudaStream_t st) {
  const int Kq = (K + 63) & ~63;
  const int tiles_per_img = cdiv(HW, TM);
  const long long grid = (long long)n_img * tiles_per_img;
  ADLM_CHECK_SUPPORTED(grid < (1ll << 31), "too many pixel tiles");
  const size_t bytes = sizeof(float) * ((size_t)Kq * TM + (size_t)Kq * PS + TM + TN) +
                       sizeof(unsigned long long) * TN + sizeof(uint32_t) * TM + sizeof(int) * TN;
  ADLM_CUDA(cudaFuncSetAttribute(k_simt_push_scan, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  KernelTimer timer(ADLM_K_PUSH_SCAN, st);
  k_simt_push_scan<<<(unsigned)grid, NT, bytes, st>>>(z, K, HW, mask, P, N, proto_class,
                                                      (unsigned long long)patch_base, packed_min, tiles_per_img,
                                                      per_image);
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

int push_gather_impl(const float* z, int n_img, int K, int HW, long long patch_base,
                     const unsigned long long* packed_min, int N, float* proto_out, cudaStream_t st) {
  k_push_gather<<<N, 64, 0, st>>>(z, n_img, K, HW, (unsigned long long)patch_base, packed_min, N, proto_out);
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

}  // namespace adlm
