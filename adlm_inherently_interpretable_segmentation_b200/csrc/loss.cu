// Loss side of the fused step: label histogram, cross entropy (+ its gradient), the KLD
// prototype-diversity loss on distances (+ its gradient), masked L1, final scalars.
//
// Reference: segmentation/module.py:143-222.  Quirks kept on purpose:
//   * KLD groups are keyed by the UN-shifted label (module.py:150,170-178): with ignore_void the
//     void pixels (label 0) pair up with class-0 prototypes, label l with class l, label C is skipped;
//   * the KLD runs on distances, not similarities (module.py:142,153);
//   * groups with < 2 prototypes or < 2 pixels are skipped (module.py:186-188);
//   * kld_loss = mean over all (image, label, pair) of exp(-(kl(p2||p1) + kl(p1||p2)) / 2).
//
// Algebra used (la = log_softmax over the group's pixels, q = exp(la), E_a[f] = sum_m q_a(m) f(m)):
//   kl_div(la_i, la_j, log_target) = sum q_j (la_j - la_i)  =>  kld_ij = (E_j[d_j-d_i] - E_i[d_j-d_i]) / 2
// so one J x J matrix A[a][b] = E_a[d_b] per group gives every pair.  Gradient wrt d_a(m):
//   sum_b c_ab * ( q_a ((d_a - d_b) - (A[a][a] - A[a][b])) + q_a - q_b ),  c_ab = -(w/T) exp(-kld_ab) / 2.
#include <atomic>

#include "common.cuh"

namespace adlm {

constexpr int JMAX = ADLM_MAX_PROTOS_PER_CLASS;
namespace {

// One thread per pixel: CE contribution, correctness, dlogits (module.py:156-165,210-211).
__global__ void k_ce(const float* __restrict__ logits, const int32_t* __restrict__ target, long long M,
                     int C, int ignore_void, float w_ce_scaled, const int* __restrict__ counts,
                     float* __restrict__ dlogits, double* __restrict__ accum) {
  const long long m = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  float loss = 0.f;
  int correct = 0;
  if (m < M) {
    const float* lg = logits + m * C;
    int t = target[m];
    bool valid;
    if (ignore_void) { valid = (t >= 1 && t <= C); t -= 1; }
    else             { valid = (t >= 0 && t < C); }
    if (valid) {
      float mx = lg[0];
      int am = 0;
      for (int c = 1; c < C; ++c) { const float v = lg[c]; if (v > mx) { mx = v; am = c; } }
      float se = 0.f;
      for (int c = 0; c < C; ++c) se += expf(lg[c] - mx);
      const float lse = mx + logf(se);
      loss = lse - lg[t];
      correct = (am == t);
      if (dlogits) {
        const float sc = w_ce_scaled / (float)counts[0];
        for (int c = 0; c < C; ++c) {
          const float pr = expf(lg[c] - lse);
          dlogits[m * C + c] = sc * (pr - (c == t ? 1.f : 0.f));
        }
      }
    } else if (dlogits) {
      for (int c = 0; c < C; ++c) dlogits[m * C + c] = 0.f;
    }
  }
  loss = warp_sum(loss);
  const float fc = warp_sum((float)correct);
  if ((threadIdx.x & 31) == 0) {
    if (loss != 0.f) atomicAdd(&accum[0], (double)loss);
    if (fc != 0.f) atomicAdd(&accum[1], (double)fc);
  }
}

// One CTA per (label l = blockIdx.x, image b = blockIdx.y): the group's pixels are first compacted into a
// shared-memory index list (a group is ~1/C of the image), then: per-prototype max, the J x J matrix
// A[a][b] = E_a[d_b], the pair terms, and — same CTA, which owns exactly these pixels — dL/d(own distances)
// for the backward kernels.  Thread = (pixel lane tid/16, prototype-in-class tid%16): ~50 registers, so all
// B*C CTAs are resident at once.
constexpr int KLD_THREADS = 512, KLD_PL = KLD_THREADS / 16, KLD_WARPS = KLD_THREADS / 32, KLD_U = 8;

// dL/d(own distances) of the pixels of one (image, label) group, see k_kld.  JJ = J rounded up to 12 or 16 keeps
// the loops unrolled; rows / columns >= J of kl are zero, the distances beyond J (never written by the forward
// kernel) are replaced by zero so that stale workspace bits cannot poison the sums.
template <int JJ>
__device__ __forceinline__ void kld_grad_pass(const float* __restrict__ ob, float* __restrict__ og,
                                              const int* __restrict__ s_list, int n, int J,
                                              const float (*s_kl)[JMAX], const float* __restrict__ s_cs,
                                              const float* __restrict__ s_r, const float* __restrict__ s_lse,
                                              int tid) {
  for (int i = tid; i < n; i += KLD_THREADS) {
    const int pidx = s_list[i];
    const float4* row = reinterpret_cast<const float4*>(ob + (size_t)pidx * JMAX);
    float d[JJ], q[JJ];
#pragma unroll
    for (int v = 0; v < JJ / 4; ++v) {
      const float4 x = row[v];
      d[4 * v] = x.x; d[4 * v + 1] = x.y; d[4 * v + 2] = x.z; d[4 * v + 3] = x.w;
    }
#pragma unroll
    for (int c = 0; c < JJ; ++c) {
      d[c] = c < J ? d[c] : 0.f;
      q[c] = c < J ? __expf(d[c] - s_lse[c]) : 0.f;
    }
    float4* dst = reinterpret_cast<float4*>(og + (size_t)pidx * JMAX);
#pragma unroll
    for (int a4 = 0; a4 < JMAX / 4; ++a4) {
      float g[4] = {0.f, 0.f, 0.f, 0.f};
      if (a4 < JJ / 4) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int a = a4 * 4 + e;
          float cd = 0.f, cq = 0.f;
#pragma unroll
          for (int v = 0; v < JJ / 4; ++v) {
            const float4 k4 = *reinterpret_cast<const float4*>(&s_kl[a][4 * v]);
            cd = fmaf(k4.x, d[4 * v], cd);         cq = fmaf(k4.x, q[4 * v], cq);
            cd = fmaf(k4.y, d[4 * v + 1], cd);     cq = fmaf(k4.y, q[4 * v + 1], cq);
            cd = fmaf(k4.z, d[4 * v + 2], cd);     cq = fmaf(k4.z, q[4 * v + 2], cq);
            cd = fmaf(k4.w, d[4 * v + 3], cd);     cq = fmaf(k4.w, q[4 * v + 3], cq);
          }
          const float csa = s_cs[a];
          g[e] = q[a] * (d[a] * csa - cd - s_r[a]) + q[a] * csa - cq;
          asm volatile("" ::: "memory");           // keep the next row's coefficient loads below this point
        }
      }
      dst[a4] = make_float4(g[0], g[1], g[2], g[3]);
    }
  }
}

__global__ void __launch_bounds__(KLD_THREADS, 2)
k_kld(const float* __restrict__ own, const int32_t* __restrict__ target, const int* __restrict__ hist,
      const int* __restrict__ cls_count, const int* __restrict__ counts, int HW, int C,
      float coef /* w_kld * grad_scale */, float* __restrict__ own_grad, double* __restrict__ accum) {
  static_assert(JMAX == 16, "thread layout assumes 16 prototypes per class at most");
  extern __shared__ int s_list[];                 // [cnt] pixel indices of this group
  // the dZ kernel may start on SMs this grid leaves idle (1.03 waves: a long tail); it waits for this grid before
  // reading the own-class gradients
  asm volatile("griddepcontrol.launch_dependents;");
  const int l = blockIdx.x, b = blockIdx.y;
  const int J = min(cls_count[l], JMAX);
  const int cnt = hist[b * (C + 1) + l];
  const int tid = threadIdx.x, lane = tid & 31, j = tid & 15, pl = tid >> 4;
  if (cnt == 0) return;
  __shared__ int s_n;
  __shared__ float s_mx[JMAX], s_S[JMAX], s_lse[JMAX], s_cs[JMAX], s_r[JMAX];
  __shared__ float s_T[JMAX][JMAX];
  __shared__ float s_red[KLD_WARPS][JMAX];
  __shared__ __align__(16) float s_kl[JMAX][JMAX];
  __shared__ float s_part[KLD_WARPS * (JMAX + 1) * JMAX];   // per-warp partial sums of pass B
  const int32_t* t = target + (size_t)b * HW;
  const float* ob = own + (size_t)b * HW * JMAX;
  // Compaction without atomics: every warp counts its hits (labels are re-read from L1 in the second sweep), a
  // 16-entry prefix gives each warp its base, and the ballots place the indices.  (One shared-memory atomicAdd
  // with a returned value per ballot — 16 warps x 12 dependent same-address atomics at h*w = 4225 — cost more
  // than the three passes over the distances together.)
  __shared__ int s_wcnt[KLD_WARPS];
  const int wid = tid >> 5;
  int mine = 0;
  for (int p0 = 0; p0 < HW; p0 += 4 * KLD_THREADS) {
    int lab[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int pp = p0 + u * KLD_THREADS + tid;
      lab[u] = pp < HW ? t[pp] : -1;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) mine += __popc(__ballot_sync(0xffffffffu, lab[u] == l));
  }
  if (lane == 0) s_wcnt[wid] = mine;              // warp-uniform
  __syncthreads();
  int base = 0, total = 0;
#pragma unroll
  for (int w = 0; w < KLD_WARPS; ++w) {
    const int c = s_wcnt[w];
    base += w < wid ? c : 0;
    total += c;
  }
  for (int p0 = 0; p0 < HW; p0 += 4 * KLD_THREADS) {
    int pp[4];
    bool hit[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      pp[u] = p0 + u * KLD_THREADS + tid;
      hit[u] = pp[u] < HW && t[pp[u]] == l;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const unsigned bal = __ballot_sync(0xffffffffu, hit[u]);
      if (hit[u]) s_list[base + __popc(bal & ((1u << lane) - 1u))] = pp[u];
      base += __popc(bal);
    }
  }
  if (tid == 0) s_n = total;
  __syncthreads();
  const int n = s_n;                              // == cnt
  const bool valid_group = J >= 2 && n >= 2;
  float* og = own_grad ? own_grad + (size_t)b * HW * JMAX : nullptr;
  if (!valid_group) {
    if (og)
      for (int i = pl; i < n; i += KLD_PL) og[(size_t)s_list[i] * JMAX + j] = 0.f;
    return;
  }
  // The three passes over the group's pixels are chains of gathers from L2; each batches KLD_U pixel iterations so
  // that a thread has KLD_U independent loads in flight instead of one round trip per iteration.
  // pass A: per-prototype max over the group's pixels
  float mx = -INFINITY;
  for (int i0 = pl; i0 < n; i0 += KLD_PL * KLD_U) {
    float v[KLD_U];
#pragma unroll
    for (int u = 0; u < KLD_U; ++u) {
      const int i = i0 + u * KLD_PL;
      v[u] = i < n ? ob[(size_t)s_list[i] * JMAX + j] : -INFINITY;
    }
#pragma unroll
    for (int u = 0; u < KLD_U; ++u) mx = fmaxf(mx, v[u]);
  }
  mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
  if (lane < 16) s_red[tid >> 5][j] = mx;
  __syncthreads();
  if (tid < JMAX) {
    float m = s_red[0][tid];
    for (int w = 1; w < KLD_WARPS; ++w) m = fmaxf(m, s_red[w][tid]);
    s_mx[tid] = m;
  }
  __syncthreads();
  // pass B: S_j = sum exp(d_j - mx_j), T[j][c] = sum exp(d_j - mx_j) d_c; per-warp partials go through
  // s_part[warp][c][j] (conflict-free, no atomics) and are summed by 16 x 17 threads
  {
    const float mxj = s_mx[j];
    float S = 0.f, T[JMAX];
#pragma unroll
    for (int c = 0; c < JMAX; ++c) T[c] = 0.f;
    constexpr int UB = 2;
    for (int i0 = pl; i0 < n; i0 += KLD_PL * UB) {
      float4 x[UB][JMAX / 4];
      float dj[UB];
#pragma unroll
      for (int u = 0; u < UB; ++u) {
        const int i = i0 + u * KLD_PL;
        const float* row = ob + (size_t)s_list[i < n ? i : 0] * JMAX;
        dj[u] = i < n ? row[j] : -INFINITY;
        const float4* o4 = reinterpret_cast<const float4*>(row);
#pragma unroll
        for (int v = 0; v < JMAX / 4; ++v) x[u][v] = o4[v];
      }
#pragma unroll
      for (int u = 0; u < UB; ++u) {
        const float e = (j < J) ? __expf(dj[u] - mxj) : 0.f;      // exp(-inf) = 0 for the padding iterations
        S += e;
#pragma unroll
        for (int v = 0; v < JMAX / 4; ++v) {
          T[4 * v] = fmaf(e, x[u][v].x, T[4 * v]);
          T[4 * v + 1] = fmaf(e, x[u][v].y, T[4 * v + 1]);
          T[4 * v + 2] = fmaf(e, x[u][v].z, T[4 * v + 2]);
          T[4 * v + 3] = fmaf(e, x[u][v].w, T[4 * v + 3]);
        }
      }
    }
    S += __shfl_xor_sync(0xffffffffu, S, 16);
#pragma unroll
    for (int c = 0; c < JMAX; ++c) T[c] += __shfl_xor_sync(0xffffffffu, T[c], 16);
    if (lane < 16) {
      float* part = s_part + (size_t)(tid >> 5) * (JMAX + 1) * JMAX;
#pragma unroll
      for (int c = 0; c < JMAX; ++c) part[c * JMAX + j] = T[c];
      part[JMAX * JMAX + j] = S;
    }
  }
  __syncthreads();
  if (tid < (JMAX + 1) * JMAX) {
    float a = 0.f;
#pragma unroll
    for (int w = 0; w < KLD_WARPS; ++w) a += s_part[(size_t)w * (JMAX + 1) * JMAX + tid];
    const int c = tid / JMAX, jj = tid % JMAX;
    if (c < JMAX) s_T[jj][c] = a; else s_S[jj] = a;
  }
  __syncthreads();
  if (tid < JMAX * JMAX) {
    const int a = tid / JMAX, i = tid % JMAX;
    if (a < J && i < J) s_T[a][i] = s_T[a][i] / s_S[a];     // A[a][i] = E_a[d_i]
  }
  __syncthreads();
  const float n_terms = (float)counts[1];
  float esum = 0.f;
  if (tid < JMAX * JMAX) {
    const int a = tid / JMAX, c = tid % JMAX;
    float cv = 0.f;
    if (a < J && c < J && a != c) {
      const float kld = 0.5f * (s_T[a][a] - s_T[a][c] - s_T[c][a] + s_T[c][c]);
      const float e = __expf(-kld);
      if (a < c) esum = e;
      cv = -0.5f * coef * e / n_terms;
    }
    s_kl[a][c] = cv;
  }
  esum = warp_sum(esum);
  if (lane == 0 && esum != 0.f) atomicAdd(&accum[2], (double)esum);
  __syncthreads();
  if (tid < JMAX) {
    const int a = tid;
    float cs = 0.f, r = 0.f;
    if (a < J) {
      for (int c = 0; c < J; ++c) {
        cs += s_kl[a][c];
        r = fmaf(s_kl[a][c], s_T[a][a] - s_T[a][c], r);
      }
    }
    s_cs[a] = cs;
    s_r[a] = r;
    s_lse[a] = (a < J) ? s_mx[a] + __logf(s_S[a]) : 0.f;
  }
  __syncthreads();
  if (og == nullptr) return;
  // gradient wrt the own-class distances: thread = pixel.  The pixel's row sits in registers and the J x J
  // coefficient matrix is read as float4 broadcasts from shared memory: 2 J^2 FMAs per pixel and no shuffles (the
  // previous layout, 16 threads per pixel exchanging d and q by 32 shuffles, executed 3.4x the instructions).
  if (J <= 12) kld_grad_pass<12>(ob, og, s_list, n, J, s_kl, s_cs, s_r, s_lse, tid);
  else         kld_grad_pass<16>(ob, og, s_list, n, J, s_kl, s_cs, s_r, s_lse, tid);
}

// Prologue of the step in one launch: one CTA per image builds that image's label histogram in shared memory
// and stores it; the last CTA to finish (ticket) derives the class sizes, the two loss normalisers and the
// prototype ranks and clears the accumulators.  The ticket counters live in a small device array (zero at
// module load, reset by the last CTA); the host rotates through the slots so that steps in flight on
// different streams never share one.
constexpr int TICKET_SLOTS = 64;
__device__ unsigned int g_prepare_ticket[TICKET_SLOTS];

__global__ void __launch_bounds__(1024)
k_step_prepare(const int32_t* __restrict__ target, const int32_t* __restrict__ proto_class, int B, int HW, int C,
               int N, int ignore_void, int slot, int* __restrict__ hist, int* __restrict__ cls_count,
               int* __restrict__ counts, int32_t* __restrict__ rank, double* __restrict__ accum,
               const DpComm comm, const DpLayout lay) {
  __shared__ int sh[ADLM_MAX_C + 1];
  __shared__ int s_cls[ADLM_MAX_C];
  __shared__ int s_valid, s_terms, s_last;
  const int tid = threadIdx.x, b = blockIdx.x;
  // the forward kernel may start now: it stages its operands and waits for this grid before reading its outputs
  asm volatile("griddepcontrol.launch_dependents;");
  // Everything that depends on the prototype classes only (class sizes for every CTA, ranks by CTA 0) runs while the
  // label loads are in flight, not in the tail of the last CTA.
  __shared__ int s_pc[1024];
  for (int i = tid; i <= C; i += 1024) sh[i] = 0;
  for (int c = tid; c < C; c += 1024) s_cls[c] = 0;
  if (tid == 0) { s_valid = 0; s_terms = 0; }
  const int32_t* t = target + (size_t)b * HW;
  constexpr int LB = 8;                           // label loads in flight per thread
  int lab[LB];
#pragma unroll
  for (int u = 0; u < LB; ++u) {
    const int p = u * 1024 + tid;
    lab[u] = p < HW ? t[p] : -1;
  }
  const int pc0 = tid < N ? proto_class[tid] : -2;
  __syncthreads();
#pragma unroll
  for (int u = 0; u < LB; ++u)
    if (lab[u] >= 0 && lab[u] <= C) atomicAdd(&sh[lab[u]], 1);
  for (int p = LB * 1024 + tid; p < HW; p += 1024) {
    const int l = t[p];
    if (l >= 0 && l <= C) atomicAdd(&sh[l], 1);
  }
  // class sizes (every CTA: the last one needs them) and the rank of a prototype inside its class (CTA 0; kept in a
  // register until the wait below: nothing is WRITTEN to global memory before it)
  int my_rank = 0;
  for (int n0 = 0; n0 < N; n0 += 1024) {
    const int n = n0 + tid;
    const int c = n0 == 0 ? pc0 : (n < N ? proto_class[n] : -2);
    if (n0 > 0) __syncthreads();
    s_pc[tid] = c;
    __syncthreads();
    if (n < N) {
      if (c >= 0 && c < C) atomicAdd(&s_cls[c], 1);
      if (b == 0 && n0 == 0) {
        int r = 0;
        for (int i = 0; i < tid; ++i) r += (s_pc[i] == c);
        my_rank = r;
      }
    }
  }
  __syncthreads();
  // Back-to-back steps: this grid may have started while the previous step's partials reduction still runs (it reads
  // the accumulators this kernel clears and shares the workspace).  Everything above only READS inputs of this step
  // (labels, prototype classes) into registers / shared memory; every global write comes after this wait.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (b == 0) {
    // a rank beyond the per-class capacity of the KLD side buffers is stored as -1: the head kernels then leave that
    // prototype out of the KLD term and the last CTA flags the step as failed (scalars[7] > 0)
    if (tid < N) rank[tid] = my_rank < JMAX ? my_rank : -1;
    for (int n = 1024 + tid; n < N; n += 1024) {
      const int c = proto_class[n];
      int r = 0;
      for (int i = 0; i < n; ++i) r += (proto_class[i] == c);
      rank[n] = r < JMAX ? r : -1;
    }
  }
  for (int i = tid; i <= C; i += 1024) hist[b * (C + 1) + i] = sh[i];
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const unsigned int prev = atomicAdd(&g_prepare_ticket[slot], 1u);
    s_last = (prev == (unsigned int)(B - 1));
    if (s_last) g_prepare_ticket[slot] = 0u;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  int valid = 0, terms = 0;
  const volatile int* vh = hist;
  for (int i = tid; i < B * (C + 1); i += 1024) {
    const int l = i % (C + 1), cnt = vh[i];
    if (ignore_void) { if (l >= 1) valid += cnt; }
    else             { if (l < C) valid += cnt; }
    if (l < C && cnt >= 2 && s_cls[l] >= 2) terms += s_cls[l] * (s_cls[l] - 1) / 2;
  }
  if (valid) atomicAdd(&s_valid, valid);
  if (terms) atomicAdd(&s_terms, terms);
  __syncthreads();
  for (int c = tid; c < C; c += 1024) cls_count[c] = s_cls[c];
  if (tid < 4) accum[tid] = 0.0;
  __syncthreads();
  // more prototypes in one class than the own-class side buffers hold: the KLD term would be wrong -> fail the step
  for (int c = tid; c < C; c += 1024)
    if (s_cls[c] > JMAX) accum[3] = 1.0;
  if (comm.seq == 0u) {
    if (tid == 0) { counts[0] = s_valid; counts[1] = s_terms; }
    return;
  }
  // data-parallel: publish this shard's normalisers to every peer, wait for theirs, sum in rank order
  __shared__ int s_all[DP_MAX_WORLD][2];
  __shared__ int s_timeout;
  if (tid == 0) s_timeout = 0;
  __syncthreads();
  const int par = (int)(comm.seq & 1u), Wd = comm.world;
  if (tid < Wd) {
    char* pb = comm.peer[tid];
    volatile int* slot_out = reinterpret_cast<volatile int*>(pb + lay.cnt) + (par * Wd + comm.rank) * 2;
    slot_out[0] = s_valid;
    slot_out[1] = s_terms;
    __threadfence_system();
    st_release_sys(reinterpret_cast<unsigned int*>(pb + lay.cflag) + par * Wd + comm.rank, comm.seq);
  }
  if (comm.defer_counts) {
    // the rendezvous is taken off this kernel's tail: the forward collects the peers' counts when its first logits
    // epilogue needs them and stores the sums over `counts` (until then it holds this shard's own)
    if (tid == 0) { counts[0] = s_valid; counts[1] = s_terms; }
    return;
  }
  if (tid < Wd) {
    char* me = comm.peer[comm.rank];
    const bool ok = dp_wait_flag(reinterpret_cast<const unsigned int*>(me + lay.cflag) + par * Wd + tid, comm.seq,
                                 comm.timeout_ns);
    const volatile int* slot_in = reinterpret_cast<const volatile int*>(me + lay.cnt) + (par * Wd + tid) * 2;
    s_all[tid][0] = slot_in[0];
    s_all[tid][1] = slot_in[1];
    if (!ok) s_timeout = 1;
  }
  __syncthreads();
  if (tid == 0) {
    int v = 0, t2 = 0;
    for (int r = 0; r < Wd; ++r) { v += s_all[r][0]; t2 += s_all[r][1]; }
    counts[0] = v;
    counts[1] = t2;
    if (s_timeout) accum[3] = 1.0;
  }
}

// l1 = |W * (1 - identity^T)|_1 (module.py:213-218), dW += w_l1 * sign(W) * mask, final scalars.
__global__ void k_finalize(const float* __restrict__ W, const int32_t* __restrict__ proto_class, int C, int N,
                           float w_ce, float w_kld, float w_l1, float grad_scale,
                           const int* __restrict__ counts, const double* __restrict__ accum,
                           float* __restrict__ dW, double* __restrict__ scalars) {
  __shared__ float s_part[32];
  float l1 = 0.f;
  for (int i = threadIdx.x; i < C * N; i += blockDim.x) {
    const int c = i / N, n = i - c * N;
    if (proto_class[n] != c) {
      const float w = W[i];
      l1 += fabsf(w);
      if (dW) dW[i] += w_l1 * grad_scale * ((w > 0.f) ? 1.f : ((w < 0.f) ? -1.f : 0.f));
    }
  }
  l1 = warp_sum(l1);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = l1;
  __syncthreads();
  if (threadIdx.x == 0) {
    double tot = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += (double)s_part[w];
    const double n_valid = (double)counts[0], n_terms = (double)counts[1];
    const double ce = n_valid > 0 ? accum[0] / n_valid
                                   : __longlong_as_double(0x7ff8000000000000LL);  // empty mean = NaN, as torch
    const double kld = n_terms > 0 ? accum[2] / n_terms : 0.0;
    scalars[0] = (double)w_ce * ce + (double)w_kld * kld + (double)w_l1 * tot;
    scalars[1] = ce;
    scalars[2] = kld;
    scalars[3] = tot;
    scalars[4] = accum[1];
    scalars[5] = n_valid;
    scalars[6] = n_terms;
    scalars[7] = accum[3];
  }
}

}  // namespace

// ---------------------------------------------------------------------------------- host side
static std::atomic<unsigned int> g_next_ticket_slot{0};
struct StepWorkspace {  // carved out of the caller's workspace buffer
  float* logits;     // [M, C]
  float* dlogits;    // [M, C]
  float* own;        // [M, JMAX]
  float* own_grad;   // [M, JMAX]
  int* hist;         // [B, C+1]
  int* cls_count;    // [C]
  int* counts;       // [2]
  int32_t* rank;     // [N]
  double* accum;     // [4]
  void* bwd_scratch; // per-CTA dP/dW partials of the tensor-core backward
  size_t bwd_scratch_bytes;
  size_t bytes;
};

StepWorkspace carve_step_workspace(void* base, int B, int K, int HW, int N, int C) {
  StepWorkspace w;
  const size_t M = (size_t)B * HW;
  size_t off = 0;
  auto take = [&](size_t bytes) { void* p = base ? (char*)base + off : nullptr; off += align_up(bytes, 256); return p; };
  w.logits = (float*)take(M * C * sizeof(float));
  w.dlogits = (float*)take(M * C * sizeof(float));
  w.own = (float*)take(M * JMAX * sizeof(float));
  w.own_grad = (float*)take(M * JMAX * sizeof(float));
  w.hist = (int*)take((size_t)B * (C + 1) * sizeof(int));
  w.cls_count = (int*)take((size_t)C * sizeof(int));
  w.counts = (int*)take(2 * sizeof(int));
  w.rank = (int32_t*)take((size_t)N * sizeof(int32_t));
  w.accum = (double*)take(4 * sizeof(double));
  w.bwd_scratch_bytes = head_bwd_workspace_bytes(ADLM_PREC_TF32, K, N);
  w.bwd_scratch = take(w.bwd_scratch_bytes);
  w.bytes = off;
  return w;
}

size_t step_workspace_bytes(int B, int K, int HW, int N, int C) {
  return carve_step_workspace(nullptr, B, K, HW, N, C).bytes;
}

int step_fwd_bwd_impl(const HeadShape& s, int prec_mode, const float* z, const float* P, const float* W,
                      const int32_t* target, const int32_t* proto_class, int ignore_void,
                      float w_ce, float w_kld, float w_l1, float grad_scale, const int32_t* global_counts,
                      float* logits_out, int32_t* argmax_out, float* dz, float* dP, float* dW,
                      double* scalars, void* workspace, const DpComm* comm, float* z_out, cudaStream_t st) {
  // z_out != null: `z` holds PRE-sigmoid features; the forward writes z = sigmoid(z) to z_out and dz is the gradient
  // wrt the pre-sigmoid features (fused add-on layer, model.py:124-128,174)
  const int B = s.B, HW = s.HW, N = s.N, C = s.C;
  const long long M = (long long)B * HW;
  StepWorkspace ws = carve_step_workspace(workspace, B, s.K, HW, N, C);
  const bool want_grad = (dz != nullptr) || (dP != nullptr) || (dW != nullptr);
  float* logits = logits_out ? logits_out : ws.logits;

  {
    const int slot = (int)(g_next_ticket_slot.fetch_add(1u, std::memory_order_relaxed) % TICKET_SLOTS);
    DpComm cm;
    DpLayout lay = {};
    if (comm != nullptr) {
      if (!(want_grad && dz && head_fwd_fuses_ce(s, prec_mode) && head_bwd_fuses_finalize(s, prec_mode, dP, dW))) {
        set_error("adlm_step_fwd_bwd_dp: needs ADLM_PREC_TF32, K <= 64, N <= 192, C <= 32 and dz, dP, dW");
        return ADLM_E_UNSUPPORTED;
      }
      cm = *comm;
      lay = dp_layout(s.K, N, C, cm.world);
    }
    KernelTimer timer(ADLM_K_STEP_PREPARE, st);
    ADLM_CUDA(launch_maybe_pdl(true, k_step_prepare, dim3(B), dim3(1024), 0, st, target, proto_class, B, HW, C, N,
                               ignore_void, slot, ws.hist, ws.cls_count, ws.counts, ws.rank, ws.accum, cm, lay));
    ADLM_LAUNCH_CHECK();
  }
  const int* counts = (global_counts && !comm) ? global_counts : ws.counts;

  StepExtras ex;
  ex.target = target;
  ex.proto_class = proto_class;
  ex.proto_rank = ws.rank;
  ex.own_dist = ws.own;
  const bool fuse_ce = head_fwd_fuses_ce(s, prec_mode);
  if (z_out != nullptr) {
    if (!head_fuses_sigmoid(s, prec_mode) || !fuse_ce) {
      set_error("fused sigmoid prologue needs ADLM_PREC_TF32, K <= 64, N <= 192, C <= 32");
      return ADLM_E_UNSUPPORTED;
    }
    ex.z_out = z_out;
  }
  if (fuse_ce) {
    // CE, accuracy and dlogits come out of the forward kernel's logits registers; logits touch HBM only on request
    ex.ce_counts = counts;
    if (comm != nullptr && comm->defer_counts) {
      ex.ce_dp = *comm;
      ex.ce_dp_lay = dp_layout(s.K, N, C, comm->world);
      ex.ce_counts_out = ws.counts;
    }
    ex.after_prepare = 1;                  // the forward is the launch right after k_step_prepare
    ex.ce_dlogits = want_grad ? ws.dlogits : nullptr;
    ex.ce_accum = ws.accum;
    ex.ce_scale = w_ce * grad_scale;
    ex.ignore_void = ignore_void;
  }
  int rc;
  if ((rc = head_fwd_dispatch(s, prec_mode, z, P, W, fuse_ce ? logits_out : logits, nullptr, nullptr, argmax_out, ex,
                              st)) != ADLM_OK)
    return rc;
  if (!fuse_ce) {
    k_ce<<<(unsigned)((M + 127) / 128), 128, 0, st>>>(logits, target, M, C, ignore_void, w_ce * grad_scale,
                                                      counts, want_grad ? ws.dlogits : nullptr, ws.accum);
    ADLM_LAUNCH_CHECK();
  }
  {
    dim3 grid(C, B);
    ADLM_CHECK_SUPPORTED((size_t)HW * sizeof(int) <= 200 * 1024, "feature map too large for the KLD kernel (h*w > 51200)");
    ADLM_CUDA(cudaFuncSetAttribute(k_kld, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)((size_t)HW * sizeof(int))));
    KernelTimer timer(ADLM_K_KLD, st);
    k_kld<<<grid, KLD_THREADS, (size_t)HW * sizeof(int), st>>>(ws.own, target, ws.hist, ws.cls_count, counts, HW, C,
                                                    w_kld * grad_scale, want_grad ? ws.own_grad : nullptr, ws.accum);
    ADLM_LAUNCH_CHECK();
  }
  bool fin_fused = false;
  if (want_grad) {
    StepExtras exb = ex;
    exb.own_dist = nullptr;
    exb.own_grad = ws.own_grad;
    exb.z_out = nullptr;
    exb.dz_sigmoid = z_out != nullptr ? 1 : 0;
    if (head_bwd_fuses_finalize(s, prec_mode, dP, dW)) {
      exb.fin.scalars = scalars;
      exb.fin.counts = counts;
      exb.fin.accum = ws.accum;
      exb.fin.w_ce = w_ce; exb.fin.w_kld = w_kld; exb.fin.w_l1 = w_l1; exb.fin.grad_scale = grad_scale;
      if (comm != nullptr) {
        exb.fin.comm = *comm;
        exb.fin.lay = dp_layout(s.K, N, C, comm->world);
      }
      fin_fused = true;
    }
    if ((rc = head_bwd_dispatch(s, prec_mode, z_out ? z_out : z, P, W, ws.dlogits, nullptr, nullptr, dz, dP, dW, exb, ws.bwd_scratch,
                                ws.bwd_scratch_bytes, st)) != ADLM_OK)
      return rc;
  }
  if (!fin_fused) {
    k_finalize<<<1, 1024, 0, st>>>(W, proto_class, C, N, w_ce, w_kld, w_l1, grad_scale, counts, ws.accum,
                                  dW, scalars);
    ADLM_LAUNCH_CHECK();
  }
  return ADLM_OK;
}

int step_counts_impl(const int32_t* target, int B, int HW, const int32_t* proto_class, int N, int C,
                     int ignore_void, int32_t* counts_out, void* workspace, cudaStream_t st) {
  // the step's own prologue kernel, with the normalisers redirected to the caller's buffer (one launch)
  StepWorkspace ws = carve_step_workspace(workspace, B, 1, HW, N, C);
  const int slot = (int)(g_next_ticket_slot.fetch_add(1u, std::memory_order_relaxed) % TICKET_SLOTS);
  k_step_prepare<<<B, 1024, 0, st>>>(target, proto_class, B, HW, C, N, ignore_void, slot, ws.hist, ws.cls_count,
                                     counts_out, ws.rank, ws.accum, DpComm{}, DpLayout{});
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// adlm_step_counts_noalloc: the loss normalisers of a batch shard WITHOUT a caller-provided workspace.  No counterpart
// in the reference; added for callers that cannot allocate (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The scratch of k_step_prepare (per-image
// label histograms, class sizes, prototype ranks, accumulators) lives in a file-scope device array of fixed size that
// every GPU loading the library holds until exit.  ALL calls share it and it keeps the last call's contents: the kernel
// overwrites every word it later reads (nothing has to start at zero; the ticket counter it uses is reset by the
// kernel itself), but two calls that overlap on the GPU (different streams) race on it.
constexpr int NOALLOC_MAX_B = ADLM_NOALLOC_MAX_B, NOALLOC_MAX_N = ADLM_NOALLOC_MAX_N;
constexpr size_t NA_HIST = 0;                                                                   // int [B][C+1]
constexpr size_t NA_CLS = NA_HIST + ((size_t)NOALLOC_MAX_B * (ADLM_MAX_C + 1) * 4 + 255) / 256 * 256;   // int [C]
constexpr size_t NA_RANK = NA_CLS + ((size_t)ADLM_MAX_C * 4 + 255) / 256 * 256;                 // int32 [N]
constexpr size_t NA_ACCUM = NA_RANK + ((size_t)NOALLOC_MAX_N * 4 + 255) / 256 * 256;            // double [4]
constexpr size_t NA_BYTES = NA_ACCUM + 256;
static_assert(NA_BYTES == ADLM_NOALLOC_SCRATCH_BYTES, "include/adlm_head.h states the size of the static scratch");
static __device__ __align__(256) unsigned char g_counts_scratch[NA_BYTES];

// device address of the scratch array, looked up once per device with that device current (in host code the
// array's NAME is only a host-side stand-in for the symbol, never a device address)
static void* g_counts_scratch_ptr[64] = {};

int step_counts_noalloc_impl(const int32_t* target, int B, int HW, const int32_t* proto_class, int N, int C,
                             int ignore_void, int32_t* counts_out, cudaStream_t st) {
  if (B > NOALLOC_MAX_B || N > NOALLOC_MAX_N || C > ADLM_MAX_C) {
    set_error("adlm_step_counts_noalloc: B=%d N=%d C=%d exceed the static scratch (B <= %d, N <= %d, C <= %d); "
              "use adlm_step_counts with a workspace", B, N, C, NOALLOC_MAX_B, NOALLOC_MAX_N, ADLM_MAX_C);
    return ADLM_E_UNSUPPORTED;
  }
  int dev = 0;
  ADLM_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    set_error("adlm_step_counts_noalloc: device index %d out of range", dev);
    return ADLM_E_UNSUPPORTED;
  }
  void* base = __atomic_load_n(&g_counts_scratch_ptr[dev], __ATOMIC_ACQUIRE);
  if (base == nullptr) {
    ADLM_CUDA(cudaGetSymbolAddress(&base, g_counts_scratch));
    __atomic_store_n(&g_counts_scratch_ptr[dev], base, __ATOMIC_RELEASE);     // same value from every thread
  }
  char* b8 = static_cast<char*>(base);
  const int slot = (int)(g_next_ticket_slot.fetch_add(1u, std::memory_order_relaxed) % TICKET_SLOTS);
  k_step_prepare<<<B, 1024, 0, st>>>(target, proto_class, B, HW, C, N, ignore_void, slot,
                                     reinterpret_cast<int*>(b8 + NA_HIST), reinterpret_cast<int*>(b8 + NA_CLS),
                                     counts_out, reinterpret_cast<int32_t*>(b8 + NA_RANK),
                                     reinterpret_cast<double*>(b8 + NA_ACCUM), DpComm{}, DpLayout{});
  ADLM_LAUNCH_CHECK();
  return ADLM_OK;
}

}  // namespace adlm
