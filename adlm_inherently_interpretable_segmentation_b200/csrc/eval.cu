// Evaluation epilogue of the reference (segmentation/eval_valid.py:171-199, eval_test.py:105-106,
// push.py:243-247) as one kernel: bilinear upsampling (align_corners=False) of the logits (C planes) and of the
// prototype distances (N planes) from the feature map to image resolution, fused with the per-pixel argmax /
// argmin and the IoU / accuracy / prototype-usage counters.  The reference materialises both upsampled stacks
// (1.6 GB per 1024x2048 image for 190 prototypes) and copies them to the host; here an output tile's source patch
// sits in shared memory and only the two label planes and a few counters leave the SM.
//
// Arithmetic follows torch's upsample_bilinear2d: value = h0 * (w0 * v00 + w1 * v01) + h1 * (w0 * v10 + w1 * v11)
// with the source index / lambda tables computed on the host in float32 exactly as torch does
// (area_pixel_compute_source_index), first extremum wins as in torch.argmax / argmin.
#include "common.cuh"

namespace adlm {

namespace {

constexpr int EV_TX = 64, EV_TY = 16, EV_THREADS = 256, EV_PPT = EV_TX * EV_TY / EV_THREADS;

__global__ void __launch_bounds__(EV_THREADS)
k_eval_upsample(const float* __restrict__ logits, const float* __restrict__ dist, int h, int w, int C, int N, int H,
                int Wimg, const int32_t* __restrict__ y0l, const float* __restrict__ lyl,
                const int32_t* __restrict__ x0l, const float* __restrict__ lxl, int PR, int PC, int F,
                uint8_t* __restrict__ pred, int16_t* __restrict__ nearest, const uint8_t* __restrict__ labels,
                const int32_t* __restrict__ proto_class, unsigned long long* __restrict__ counters) {
  extern __shared__ float patch[];                     // [PR][PC][F]: F = C logits then N distances (odd)
  int* hist = reinterpret_cast<int*>(patch + (size_t)PR * PC * F);   // [2 + 2C + N] per-CTA counters
  const int tid = threadIdx.x, img = blockIdx.z;
  const int X0 = blockIdx.x * EV_TX, Y0 = blockIdx.y * EV_TY;
  const int X1 = min(X0 + EV_TX, Wimg) - 1, Y1 = min(Y0 + EV_TY, H) - 1;
  const int ys = y0l[Y0], ye = min(y0l[Y1] + 1, h - 1);
  const int xs = x0l[X0], xe = min(x0l[X1] + 1, w - 1);
  const int nr = ye - ys + 1, nc = xe - xs + 1;        // <= PR, PC (host guarantees)
  const int n_hist = 2 + 2 * C + N;
  if (counters != nullptr)
    for (int i = tid; i < n_hist; i += EV_THREADS) hist[i] = 0;
  // ---- stage the source patch: logits are pixel-major [B,h,w,C], distances plane-major [B,N,h,w]
  {
    const float* lg = logits + (size_t)img * h * w * C;
    const int tot_l = nr * nc * C;
    for (int i = tid; i < tot_l; i += EV_THREADS) {
      const int f = i % C, rc = i / C, c = rc % nc, r = rc / nc;
      patch[((size_t)r * PC + c) * F + f] = __ldg(lg + ((size_t)(ys + r) * w + xs + c) * C + f);
    }
    if (dist != nullptr) {
      const float* dm = dist + (size_t)img * N * h * w;
      const int tot_d = N * nr * nc;
      for (int i = tid; i < tot_d; i += EV_THREADS) {
        const int c = i % nc, fr = i / nc, r = fr % nr, f = fr / nr;
        patch[((size_t)r * PC + c) * F + C + f] = __ldg(dm + ((size_t)f * h + ys + r) * w + xs + c);
      }
    }
  }
  __syncthreads();
#pragma unroll 1
  for (int q = 0; q < EV_PPT; ++q) {
    const int id = tid + q * EV_THREADS;
    const int x = X0 + (id % EV_TX), y = Y0 + (id / EV_TX);
    if (x >= Wimg || y >= H) continue;
    const int yy0 = y0l[y], yy1 = min(yy0 + 1, h - 1), xx0 = x0l[x], xx1 = min(xx0 + 1, w - 1);
    const float h1 = lyl[y], h0 = 1.f - h1, w1 = lxl[x], w0 = 1.f - w1;
    const float* p00 = patch + ((size_t)(yy0 - ys) * PC + (xx0 - xs)) * F;
    const float* p01 = patch + ((size_t)(yy0 - ys) * PC + (xx1 - xs)) * F;
    const float* p10 = patch + ((size_t)(yy1 - ys) * PC + (xx0 - xs)) * F;
    const float* p11 = patch + ((size_t)(yy1 - ys) * PC + (xx1 - xs)) * F;
    float best = -INFINITY;
    int cls = 0;
#pragma unroll 4
    for (int f = 0; f < C; ++f) {
      const float v = h0 * (w0 * p00[f] + w1 * p01[f]) + h1 * (w0 * p10[f] + w1 * p11[f]);
      if (v > best || f == 0) { best = v; cls = f; }
    }
    int near = -1;
    if (dist != nullptr) {
      float bd = INFINITY;
      near = 0;
#pragma unroll 4
      for (int f = 0; f < N; ++f) {
        const float v = h0 * (w0 * p00[C + f] + w1 * p01[C + f]) + h1 * (w0 * p10[C + f] + w1 * p11[C + f]);
        if (v < bd || f == 0) { bd = v; near = f; }
      }
    }
    const size_t o = ((size_t)img * H + y) * Wimg + x;
    if (pred != nullptr) pred[o] = (uint8_t)cls;
    if (nearest != nullptr && near >= 0) nearest[o] = (int16_t)near;
    if (counters != nullptr) {
      if (labels != nullptr) {
        const int ann = labels[o];                     // 0 = void, class + 1 otherwise (eval_valid.py:178-189)
        if (ann != 0) {
          atomicAdd(&hist[1], 1);
          atomicAdd(&hist[2 + C + cls], 1);            // union: predicted class ...
          if (ann - 1 == cls) {
            atomicAdd(&hist[0], 1);
            atomicAdd(&hist[2 + cls], 1);              // intersection
          } else if (ann - 1 < C) {
            atomicAdd(&hist[2 + C + ann - 1], 1);      // ... or ground-truth class
          }
        }
      }
      // prototype usage: the nearest prototype belongs to the predicted class (eval_valid.py:192-197)
      if (near >= 0 && proto_class != nullptr && proto_class[near] == cls) atomicAdd(&hist[2 + 2 * C + near], 1);
    }
  }
  if (counters != nullptr) {
    __syncthreads();
    for (int i = tid; i < n_hist; i += EV_THREADS)
      if (hist[i] != 0) atomicAdd(&counters[i], (unsigned long long)hist[i]);
  }
}

// Upsampling factors of at least 4 in x (every segmentation config: 8): a thread owns a strip of four consecutive
// output pixels of one row.  Their left taps are column sc or sc + 1 (sc = left tap of the first pixel), so the
// vertical interpolation t_c = h0 v[y0][c] + h1 v[y1][c] is computed once for columns sc, sc+1, sc+2 and shared by
// the strip: 6 shared loads and 6 + 4 * 2 FP operations per plane for four pixels instead of 16 and 24.  The
// association differs from torch's (vertical first), which only matters for exact ties of the interpolated values.
__global__ void __launch_bounds__(EV_THREADS)
k_eval_upsample_strip(const float* __restrict__ logits, const float* __restrict__ dist, int h, int w, int C, int N, int H,
                      int Wimg, const int32_t* __restrict__ y0l, const float* __restrict__ lyl,
                      const int32_t* __restrict__ x0l, const float* __restrict__ lxl, int PR, int PC, int F,
                      uint8_t* __restrict__ pred, int16_t* __restrict__ nearest, const uint8_t* __restrict__ labels,
                      const int32_t* __restrict__ proto_class, unsigned long long* __restrict__ counters) {
  extern __shared__ float patch[];                     // [PR][PC][F]
  int* hist = reinterpret_cast<int*>(patch + (size_t)PR * PC * F);
  const int tid = threadIdx.x, img = blockIdx.z;
  const int X0 = blockIdx.x * EV_TX, Y0 = blockIdx.y * EV_TY;
  const int X1 = min(X0 + EV_TX, Wimg) - 1, Y1 = min(Y0 + EV_TY, H) - 1;
  const int ys = y0l[Y0], ye = min(y0l[Y1] + 1, h - 1);
  const int xs = x0l[X0], xe = min(x0l[X1] + 1, w - 1);
  const int nr = ye - ys + 1, nc = xe - xs + 1;
  const int n_hist = 2 + 2 * C + N;
  if (counters != nullptr)
    for (int i = tid; i < n_hist; i += EV_THREADS) hist[i] = 0;
  {
    const float* lg = logits + (size_t)img * h * w * C;
    const int tot_l = nr * nc * C;
    for (int i = tid; i < tot_l; i += EV_THREADS) {
      const int f = i % C, rc = i / C, c = rc % nc, r = rc / nc;
      patch[((size_t)r * PC + c) * F + f] = __ldg(lg + ((size_t)(ys + r) * w + xs + c) * C + f);
    }
    if (dist != nullptr) {
      const float* dm = dist + (size_t)img * N * h * w;
      const int tot_d = N * nr * nc;
      for (int i = tid; i < tot_d; i += EV_THREADS) {
        const int c = i % nc, fr = i / nc, r = fr % nr, f = fr / nr;
        patch[((size_t)r * PC + c) * F + C + f] = __ldg(dm + ((size_t)f * h + ys + r) * w + xs + c);
      }
    }
  }
  __syncthreads();
  // thread -> (row, strip): 16 strips of 4 pixels per tile row, 16 rows: one strip per thread
  const int y = Y0 + tid / (EV_TX / 4), xb = X0 + (tid % (EV_TX / 4)) * 4;
  if (y < H && xb < Wimg) {
    const int yy0 = y0l[y], yy1 = min(yy0 + 1, h - 1);
    const float h1 = lyl[y], h0 = 1.f - h1;
    const int sc = x0l[xb];
    const int c0 = sc - xs, c1 = min(sc + 1, w - 1) - xs, c2 = min(sc + 2, w - 1) - xs;
    const float* r0 = patch + (size_t)(yy0 - ys) * PC * F;
    const float* r1 = patch + (size_t)(yy1 - ys) * PC * F;
    const float *a0 = r0 + (size_t)c0 * F, *a1 = r0 + (size_t)c1 * F, *a2 = r0 + (size_t)c2 * F;
    const float *b0 = r1 + (size_t)c0 * F, *b1 = r1 + (size_t)c1 * F, *b2 = r1 + (size_t)c2 * F;
    float w0[4], w1[4];
    bool sh[4], in[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int x = min(xb + i, Wimg - 1);
      in[i] = xb + i < Wimg;
      sh[i] = x0l[x] != sc;                            // left tap is sc + 1
      w1[i] = lxl[x];
      w0[i] = 1.f - w1[i];
    }
    float best[4], bd[4];
    int cls[4], near[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { best[i] = -INFINITY; bd[i] = INFINITY; cls[i] = 0; near[i] = dist != nullptr ? 0 : -1; }
#pragma unroll 2
    for (int f = 0; f < C; ++f) {
      const float t0 = h0 * a0[f] + h1 * b0[f], t1 = h0 * a1[f] + h1 * b1[f], t2 = h0 * a2[f] + h1 * b2[f];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float v = w0[i] * (sh[i] ? t1 : t0) + w1[i] * (sh[i] ? t2 : t1);
        if (v > best[i] || f == 0) { best[i] = v; cls[i] = f; }
      }
    }
    if (dist != nullptr) {
#pragma unroll 2
      for (int f = 0; f < N; ++f) {
        const float t0 = h0 * a0[C + f] + h1 * b0[C + f], t1 = h0 * a1[C + f] + h1 * b1[C + f],
                    t2 = h0 * a2[C + f] + h1 * b2[C + f];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float v = w0[i] * (sh[i] ? t1 : t0) + w1[i] * (sh[i] ? t2 : t1);
          if (v < bd[i] || f == 0) { bd[i] = v; near[i] = f; }
        }
      }
    }
    const size_t o = ((size_t)img * H + y) * Wimg + xb;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (!in[i]) continue;
      if (pred != nullptr) pred[o + i] = (uint8_t)cls[i];
      if (nearest != nullptr && near[i] >= 0) nearest[o + i] = (int16_t)near[i];
      if (counters != nullptr) {
        if (labels != nullptr) {
          const int ann = labels[o + i];
          if (ann != 0) {
            atomicAdd(&hist[1], 1);
            atomicAdd(&hist[2 + C + cls[i]], 1);
            if (ann - 1 == cls[i]) {
              atomicAdd(&hist[0], 1);
              atomicAdd(&hist[2 + cls[i]], 1);
            } else if (ann - 1 < C) {
              atomicAdd(&hist[2 + C + ann - 1], 1);
            }
          }
        }
        if (near[i] >= 0 && proto_class != nullptr && proto_class[near[i]] == cls[i]) atomicAdd(&hist[2 + 2 * C + near[i]], 1);
      }
    }
  }
  if (counters != nullptr) {
    __syncthreads();
    for (int i = tid; i < n_hist; i += EV_THREADS)
      if (hist[i] != 0) atomicAdd(&counters[i], (unsigned long long)hist[i]);
  }
}

}  // namespace

// largest number of source rows / columns an output tile can touch, from the tables themselves
int eval_upsample_impl(const float* logits, const float* dist, int B, int h, int w, int C, int N, int H, int Wimg,
                       const int32_t* y0l, const float* lyl, const int32_t* x0l, const float* lxl, int patch_rows,
                       int patch_cols, uint8_t* pred, int16_t* nearest, const uint8_t* labels,
                       const int32_t* proto_class, unsigned long long* counters, cudaStream_t st) {
  const int nF = C + (dist != nullptr ? N : 0);
  const int F = nF | 1;                                // odd pitch: taps of neighbouring columns hit distinct banks
  const size_t bytes = (size_t)patch_rows * patch_cols * F * sizeof(float) + (size_t)(2 + 2 * C + N) * sizeof(int);
  ADLM_CHECK_SUPPORTED(bytes <= 200 * 1024, "source patch of one output tile exceeds shared memory (C + N too large)");
  ADLM_CUDA(cudaFuncSetAttribute(k_eval_upsample, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  dim3 grid((Wimg + EV_TX - 1) / EV_TX, (H + EV_TY - 1) / EV_TY, B);
  if (Wimg >= 4 * w) {
    // four consecutive output pixels span less than one source column: strip kernel
    ADLM_CUDA(cudaFuncSetAttribute(k_eval_upsample_strip, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    k_eval_upsample_strip<<<grid, EV_THREADS, bytes, st>>>(logits, dist, h, w, C, N, H, Wimg, y0l, lyl, x0l, lxl,
                                                         patch_rows, patch_cols, F, pred, nearest, labels,
                                                         proto_class, counters);
  } else {
    k_eval_upsample<<<grid, EV_THREADS, bytes, st>>>(logits, dist, h, w, C, N, H, Wimg, y0l, lyl, x0l, lxl, patch_rows,
                                                     patch_cols, F, pred, nearest, labels, proto_class, counters);
  }
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

}  // namespace adlm
