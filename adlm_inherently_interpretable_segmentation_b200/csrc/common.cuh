// Shared helpers for the ProtoSeg prototype-head kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/adlm_head.h"

namespace adlm {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);
bool sync_launches();

#define ADLM_CHECK_ARG(cond, msg)                                   \
  do {                                                              \
    if (!(cond)) {                                                  \
      adlm::set_error("%s: %s", __func__, msg);                     \
      return ADLM_E_INVALID;                                        \
    }                                                               \
  } while (0)

#define ADLM_CHECK_SUPPORTED(cond, msg)                             \
  do {                                                              \
    if (!(cond)) {                                                  \
      adlm::set_error("%s: unsupported: %s", __func__, msg);        \
      return ADLM_E_UNSUPPORTED;                                    \
    }                                                               \
  } while (0)

#define ADLM_CUDA(call)                                                              \
  do {                                                                               \
    cudaError_t e__ = (call);                                                        \
    if (e__ != cudaSuccess) {                                                        \
      adlm::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call,                   \
                      cudaGetErrorString(e__));                                      \
      return ADLM_E_CUDA;                                                            \
    }                                                                                \
  } while (0)

// ADLM_SYNC_LAUNCHES=1 in the environment makes every launch synchronous, so that a device fault is reported
// with the file:line of the kernel that raised it (debugging aid; never set for measurements)
#define ADLM_LAUNCH_CHECK()                                                          \
  do {                                                                               \
    adlm::count_launch();                                                            \
    cudaError_t e__ = cudaGetLastError();                                            \
    if (e__ == cudaSuccess && adlm::sync_launches()) e__ = cudaDeviceSynchronize();  \
    if (e__ != cudaSuccess) {                                                        \
      adlm::set_error("%s:%d launch -> %s", __FILE__, __LINE__,                      \
                      cudaGetErrorString(e__));                                      \
      return ADLM_E_CUDA;                                                            \
    }                                                                                \
  } while (0)

// RAII bracket used by the launchers: records events around the launch when bench.py asked for this tag
struct KernelTimer {
  KernelTimer(int tag, cudaStream_t st);
  ~KernelTimer();
  int slot_;
  cudaStream_t st_;
};

// Launch with programmatic stream serialisation: the grid may start while the previous kernel of the stream still
// runs (after that kernel's griddepcontrol.launch_dependents); it must execute griddepcontrol.wait before touching
// anything the previous kernel writes.  pdl == false is an ordinary launch.
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_maybe_pdl(bool pdl, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                           cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<Args&&>(args)...);
}

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// model.py:233 / :235 and their derivatives wrt d
__device__ __forceinline__ float act_fwd(float d, float eps, int act_mode) {
  if (act_mode == ADLM_ACT_LOG) return __logf(__fdividef(d + 1.0f, d + eps));
  return -d;
}
__device__ __forceinline__ float act_fwd_precise(float d, float eps, int act_mode) {
  if (act_mode == ADLM_ACT_LOG) return logf((d + 1.0f) / (d + eps));
  return -d;
}
__device__ __forceinline__ float act_grad(float d, float eps, int act_mode) {
  if (act_mode == ADLM_ACT_LOG) return 1.0f / (d + 1.0f) - 1.0f / (d + eps);
  return -1.0f;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---- launchers implemented in the .cu files -------------------------------------------------
struct HeadShape {
  int B, K, HW, N, C;
  float eps;
  int act_mode;
};

// ---- data-parallel exchange over peer-mapped memory (NVLink / NVSwitch) -----------------------
// Every rank owns one exchange buffer that all ranks of the node have mapped (adlm_dp_comm.peers).  A rank
// publishes by storing into slot [parity][its rank] of EVERY peer's buffer, then release-storing the step's
// sequence number into the matching flag; a consumer spins (acquire) on the flags in its OWN buffer and sums
// the slots in rank order, so every rank obtains bit-identical totals.  Two parities: a rank can be at most
// one step ahead of a peer (it needs the peer's counts of step s+1 to get past the prologue of step s+1, and
// those are published after the peer's step s has drained), so slot parity s&1 is never overwritten early.
constexpr int DP_MAX_WORLD = ADLM_DP_MAX_WORLD;
struct DpComm {
  int rank = 0, world = 1;
  unsigned int seq = 0;                  // 0 = exchange disabled
  int defer_counts = 0;                  // k_step_prepare publishes its counts and does NOT wait for the peers': the
                                         // forward's logits epilogue collects them at first use (tc_head_ws.cu)
  unsigned long long timeout_ns = 30000000000ull;   // bound of every wait for a peer
  char* peer[DP_MAX_WORLD] = {};
};
struct DpLayout {                        // byte offsets inside one rank's exchange buffer
  size_t cnt, cflag, acc, gflag, grad, total;
  int n_grad, n_grad_pad, n_blocks;
};
__host__ __device__ inline int dp_reduce_blocks(int K, int N, int C) { return ((N * K + C * N) * 4 + 255) / 256; }
__host__ __device__ inline DpLayout dp_layout(int K, int N, int C, int world) {
  DpLayout l;
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  l.n_grad = N * K + C * N;
  l.n_grad_pad = (l.n_grad + 63) / 64 * 64;
  l.n_blocks = dp_reduce_blocks(K, N, C);
  size_t off = 0;
  l.cnt = off;   off += up((size_t)2 * world * 2 * sizeof(int));            // [2][W][2] int32 loss normalisers
  l.cflag = off; off += up((size_t)2 * world * sizeof(unsigned int));      // [2][W] u32
  l.acc = off;   off += up((size_t)2 * world * 4 * sizeof(double));        // [2][W][4] f64 CE / correct / KLD sums
  l.gflag = off; off += up((size_t)2 * world * l.n_blocks * sizeof(unsigned int));  // [2][W][n_blocks] u32
  l.grad = off;  off += up((size_t)2 * world * l.n_grad_pad * sizeof(float));      // [2][W][n_grad_pad] f32
  l.total = off;
  return l;
}
#ifdef __CUDACC__
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// bounded spin on wall-clock nanoseconds (%globaltimer): a dead peer turns into NaN results instead of a hung GPU
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ bool dp_wait_flag(const unsigned int* p, unsigned int v, unsigned long long timeout_ns) {
  if (ld_acquire_sys(p) == v) return true;
  const unsigned long long t0 = global_ns();
  while (ld_acquire_sys(p) != v) {
    if (global_ns() - t0 > timeout_ns) return false;
    __nanosleep(40);
  }
  return true;
}
#endif

// optional extras of the fused step
struct StepFinalize {
  double* scalars = nullptr;             // [8] out; null = not requested
  const int32_t* counts = nullptr;       // [2] loss normalisers
  const double* accum = nullptr;         // [4] CE sum, correct, KLD sum, exchange time-outs
  float w_ce = 0.f, w_kld = 0.f, w_l1 = 0.f, grad_scale = 1.f;
  DpComm comm;                           // comm.seq != 0: dP/dW/sums are exchanged with the peers in the same kernel
  DpLayout lay = {};
};

struct StepExtras {
  const int32_t* target = nullptr;       // [B,HW] labels (un-shifted)
  const int32_t* proto_class = nullptr;  // [N]
  const int32_t* proto_rank = nullptr;   // [N] rank of the prototype inside its class
  float* own_dist = nullptr;             // [B*HW, JMAX] fwd: distances to the label's own class
  const float* own_grad = nullptr;       // [B*HW, JMAX] bwd: dL/d(own_dist) from the KLD term
  // fused cross entropy in the forward (module.py:156-165,210-211); honoured by the warp-specialised tcgen05
  // forward only — callers must check head_fwd_fuses_ce() and run the separate CE kernel otherwise
  const int32_t* ce_counts = nullptr;    // [0] = number of CE patches
  // data-parallel step with the loss-normaliser rendezvous deferred into the forward: the peers' counts are read
  // from this rank's exchange buffer at the first logits epilogue; CTA 0 stores the sums to ce_counts_out for the
  // kernels that follow (KLD, finalize)
  DpComm ce_dp;                          // seq != 0: deferred rendezvous
  DpLayout ce_dp_lay = {};
  int32_t* ce_counts_out = nullptr;      // [2]
  float* ce_dlogits = nullptr;           // [B*HW, C] | null
  double* ce_accum = nullptr;            // [0] += loss sum, [1] += correct
  float ce_scale = 0.f;                  // w_ce * grad_scale
  int ignore_void = 0;
  // fused add-on activation (model.py:124-128,174: add_on_layers = Sigmoid): the forward reads PRE-sigmoid
  // features, applies the sigmoid while staging the tile and writes z to z_out; the dZ kernel multiplies by
  // z (1 - z).  Honoured by the tensor-core kernels only — callers must check head_fuses_sigmoid()
  float* z_out = nullptr;                // fwd: [B,K,HW] sigmoid(input); null = input already is z
  int after_prepare = 0;                 // fwd: launched right after k_step_prepare; may overlap it (tc_head_ws.cu)
  int dz_sigmoid = 0;                    // bwd: dz <- dz * z (1 - z)
  // fused step epilogue (L1 term of dW, L1 value, the 8 scalars; module.py:213-222) inside the partials
  // reduction of the tensor-core backward — callers must check head_bwd_fuses_finalize()
  StepFinalize fin;
};
// true when head_fwd_dispatch(prec_mode, shape) runs the kernel that honours the ce_* extras
bool head_fwd_fuses_ce(const HeadShape& s, int prec_mode);
bool head_bwd_fuses_finalize(const HeadShape& s, int prec_mode, const float* dP, const float* dW);
bool head_fuses_sigmoid(const HeadShape& s, int prec_mode);

int simt_head_fwd(const HeadShape& s, const float* z, const float* P, const float* W,
                  float* logits, float* dist, float* sim, int32_t* argmax,
                  const StepExtras& ex, cudaStream_t st);
int simt_head_bwd(const HeadShape& s, const float* z, const float* P, const float* W,
                  const float* dlogits, const float* ddist, const float* dsim,
                  float* dz, float* dP, float* dW, const StepExtras& ex, cudaStream_t st);

// large heads (tc_head_big.cu): N <= 1024, K <= 256, C <= 112 on the tensor cores
bool tc_head_big_supported(const HeadShape& s);
size_t tc_head_big_workspace_bytes(int K, int N, int C);
int tc_head_fwd_big(const HeadShape& s, const float* z, const float* P, const float* W, float* logits, float* dist,
                    int32_t* argmax, void* workspace, size_t workspace_bytes, cudaStream_t st);

// precision-mode dispatch (abi.cu)
int head_fwd_dispatch(const HeadShape& s, int prec_mode, const float* z, const float* P, const float* W,
                      float* logits, float* dist, float* sim, int32_t* argmax, const StepExtras& ex,
                      cudaStream_t st);
int head_bwd_dispatch(const HeadShape& s, int prec_mode, const float* z, const float* P, const float* W,
                      const float* dlogits, const float* ddist, const float* dsim, float* dz, float* dP,
                      float* dW, const StepExtras& ex, void* workspace, size_t workspace_bytes, cudaStream_t st);
size_t head_bwd_workspace_bytes(int prec_mode, int K, int N);

// fused step (loss.cu)
struct StepWorkspace;
size_t step_workspace_bytes(int B, int K, int HW, int N, int C);
int step_fwd_bwd_impl(const HeadShape& s, int prec_mode, const float* z, const float* P, const float* W,
                      const int32_t* target, const int32_t* proto_class, int ignore_void,
                      float w_ce, float w_kld, float w_l1, float grad_scale, const int32_t* global_counts,
                      float* logits_out, int32_t* argmax_out, float* dz, float* dP, float* dW,
                      double* scalars, void* workspace, const DpComm* comm, float* z_out, cudaStream_t st);
int step_counts_impl(const int32_t* target, int B, int HW, const int32_t* proto_class, int N, int C,
                     int ignore_void, int32_t* counts_out, void* workspace, cudaStream_t st);
int step_counts_noalloc_impl(const int32_t* target, int B, int HW, const int32_t* proto_class, int N, int C,
                             int ignore_void, int32_t* counts_out, cudaStream_t st);

// push (push.cu)
int patch_class_mask_impl(const uint8_t* labels, int n_img, int H, int Wimg, const int32_t* row_lut,
                          const int32_t* col_lut, int h, int w, uint32_t* mask_out, cudaStream_t st);
int simt_push_scan(const float* z, int n_img, int K, int HW, const uint32_t* mask, const float* P, int N,
                   const int32_t* proto_class, long long patch_base, unsigned long long* packed_min,
                   int per_image, cudaStream_t st);
int min_pool_unpack_impl(const unsigned long long* packed, int B, int N, float* min_dist, int32_t* argmin,
                         cudaStream_t st);
int min_pool_bwd_impl(const float* z, int B, int K, int HW, const float* P, int N,
                      const unsigned long long* packed, const float* gmin, float* dz, float* dP, cudaStream_t st);
int push_gather_impl(const float* z, int n_img, int K, int HW, long long patch_base,
                     const unsigned long long* packed_min, int N, float* proto_out, cudaStream_t st);

int eval_upsample_impl(const float* logits, const float* dist, int B, int h, int w, int C, int N, int H, int Wimg,
                       const int32_t* y0l, const float* lyl, const int32_t* x0l, const float* lxl, int patch_rows,
                       int patch_cols, uint8_t* pred, int16_t* nearest, const uint8_t* labels,
                       const int32_t* proto_class, unsigned long long* counters, cudaStream_t st);

}  // namespace adlm
