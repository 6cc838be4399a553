// C ABI of libadlm_head.so (see include/adlm_head.h): argument checking, precision-mode dispatch,
// error reporting.  No allocation, no synchronisation: every entry point only enqueues work on the
// caller's stream.
#include <atomic>
#include <cstdarg>
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "tc_head.cuh"

namespace adlm {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
bool sync_launches() {
  static const bool on = [] { const char* v = getenv("ADLM_SYNC_LAUNCHES"); return v != nullptr && v[0] == '1'; }();
  return on;
}
void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

// ---- per-kernel event timing (bench.py roofline leg) ----
static int g_time_tag = 0, g_time_cap = 0, g_time_n = 0;
static cudaEvent_t* g_time_ev = nullptr;   // 2 per sample

KernelTimer::KernelTimer(int tag, cudaStream_t st) : slot_(-1), st_(st) {
  if (tag == g_time_tag && g_time_ev != nullptr && g_time_n < g_time_cap) {
    slot_ = g_time_n++;
    cudaEventRecord(g_time_ev[2 * slot_], st_);
  }
}
KernelTimer::~KernelTimer() {
  if (slot_ >= 0) cudaEventRecord(g_time_ev[2 * slot_ + 1], st_);
}

static int check_shape(const char* fn, int B, int K, int HW, int N, int C) {
  if (B <= 0 || K <= 0 || HW <= 0 || N <= 0 || C <= 0) {
    set_error("%s: non-positive size (B=%d K=%d HW=%d N=%d C=%d)", fn, B, K, HW, N, C);
    return ADLM_E_INVALID;
  }
  if (K > ADLM_MAX_K) { set_error("%s: unsupported: K=%d > %d", fn, K, ADLM_MAX_K); return ADLM_E_UNSUPPORTED; }
  if (C > ADLM_MAX_C) { set_error("%s: unsupported: C=%d > %d", fn, C, ADLM_MAX_C); return ADLM_E_UNSUPPORTED; }
  if ((long long)B * HW >= (1ll << 31) - 256) { set_error("%s: unsupported: B*HW >= 2^31 - 256", fn); return ADLM_E_UNSUPPORTED; }
  return ADLM_OK;
}

int head_fwd_dispatch(const HeadShape& s, int prec_mode, const float* z, const float* P, const float* W,
                      float* logits, float* dist, float* sim, int32_t* argmax, const StepExtras& ex,
                      cudaStream_t st) {
  switch (prec_mode) {
    case ADLM_PREC_FP32:
      return simt_head_fwd(s, z, P, W, logits, dist, sim, argmax, ex, st);
    case ADLM_PREC_TF32:
    case ADLM_PREC_TF32X3:
      return tc_head_fwd(s, prec_mode, z, P, W, logits, dist, sim, argmax, ex, st);
    default:
      set_error("unknown prec_mode %d", prec_mode);
      return ADLM_E_INVALID;
  }
}

int head_bwd_dispatch(const HeadShape& s, int prec_mode, const float* z, const float* P, const float* W,
                      const float* dlogits, const float* ddist, const float* dsim, float* dz, float* dP,
                      float* dW, const StepExtras& ex, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  switch (prec_mode) {
    case ADLM_PREC_FP32:
      return simt_head_bwd(s, z, P, W, dlogits, ddist, dsim, dz, dP, dW, ex, st);
    case ADLM_PREC_TF32:
    case ADLM_PREC_TF32X3:
      return tc_head_bwd(s, prec_mode, z, P, W, dlogits, ddist, dsim, dz, dP, dW, ex, workspace, workspace_bytes, st);
    default:
      set_error("unknown prec_mode %d", prec_mode);
      return ADLM_E_INVALID;
  }
}

bool head_fwd_fuses_ce(const HeadShape& s, int prec_mode) {
  return prec_mode == ADLM_PREC_TF32 && s.K <= 64 && s.N <= 224 && s.C <= 32;
}

bool head_bwd_fuses_finalize(const HeadShape& s, int prec_mode, const float* dP, const float* dW) {
  return prec_mode == ADLM_PREC_TF32 && s.K <= 64 && s.N <= 192 && s.C <= 32 && dW != nullptr && dP != nullptr;
}

bool head_fuses_sigmoid(const HeadShape& s, int prec_mode) {
  return prec_mode == ADLM_PREC_TF32 && s.K <= 64 && s.N <= 192 && s.C <= 32;
}

size_t head_bwd_workspace_bytes(int prec_mode, int K, int N) {
  if (prec_mode != ADLM_PREC_TF32 || K > 64 || N > 192) return 0;
  return tc_bwd_workspace_bytes(K, N);
}

}  // namespace adlm

using namespace adlm;

extern "C" {

const char* adlm_last_error(void) { return g_err; }
int adlm_abi_version(void) { return ADLM_ABI_VERSION; }
unsigned long long adlm_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int adlm_timing_begin(int kernel_tag, int max_samples) {
  ADLM_CHECK_ARG(kernel_tag > 0 && max_samples > 0 && max_samples <= (1 << 20), "bad tag or sample count");
  ADLM_CHECK_ARG(g_time_ev == nullptr, "timing already active");
  g_time_ev = new cudaEvent_t[2 * (size_t)max_samples];
  for (int i = 0; i < 2 * max_samples; ++i) ADLM_CUDA(cudaEventCreate(&g_time_ev[i]));
  g_time_cap = max_samples;
  g_time_n = 0;
  g_time_tag = kernel_tag;
  return ADLM_OK;
}

int adlm_timing_end(float* ms_out) {
  ADLM_CHECK_ARG(g_time_ev != nullptr && ms_out != nullptr, "timing not active or null output");
  const int n = g_time_n;
  g_time_tag = 0;
  int rc = n;
  for (int i = 0; i < n; ++i) {
    cudaError_t e = cudaEventSynchronize(g_time_ev[2 * i + 1]);
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms_out[i], g_time_ev[2 * i], g_time_ev[2 * i + 1]);
    if (e != cudaSuccess) { set_error("adlm_timing_end: %s", cudaGetErrorString(e)); rc = ADLM_E_CUDA; break; }
  }
  for (int i = 0; i < 2 * g_time_cap; ++i) cudaEventDestroy(g_time_ev[i]);
  delete[] g_time_ev;
  g_time_ev = nullptr;
  g_time_cap = g_time_n = 0;
  return rc;
}

int adlm_head_fwd(const float* z, int B, int K, int HW, const float* P, const float* W, int N, int C,
                  float eps, int act_mode, int prec_mode, float* logits, float* dist, float* sim,
                  int32_t* argmax, void* stream) {
  int rc = check_shape(__func__, B, K, HW, N, C);
  if (rc != ADLM_OK) return rc;
  ADLM_CHECK_ARG(z && P && W && logits, "z, P, W and logits must be non-null");
  ADLM_CHECK_ARG(act_mode == ADLM_ACT_LOG || act_mode == ADLM_ACT_LINEAR, "unknown act_mode");
  HeadShape s{B, K, HW, N, C, eps, act_mode};
  return head_fwd_dispatch(s, prec_mode, z, P, W, logits, dist, sim, argmax, StepExtras{}, (cudaStream_t)stream);
}

int adlm_head_bwd(const float* z, int B, int K, int HW, const float* P, const float* W, int N, int C,
                  float eps, int act_mode, int prec_mode, const float* dlogits, const float* ddist,
                  const float* dsim, float* dz, float* dP, float* dW, void* workspace, size_t workspace_bytes,
                  void* stream) {
  int rc = check_shape(__func__, B, K, HW, N, C);
  if (rc != ADLM_OK) return rc;
  ADLM_CHECK_ARG(z && P && W, "z, P and W must be non-null");
  ADLM_CHECK_ARG(dlogits || ddist || dsim, "at least one upstream gradient is required");
  ADLM_CHECK_ARG(dz || dP || dW, "at least one output gradient is required");
  ADLM_CHECK_ARG(act_mode == ADLM_ACT_LOG || act_mode == ADLM_ACT_LINEAR, "unknown act_mode");
  HeadShape s{B, K, HW, N, C, eps, act_mode};
  return head_bwd_dispatch(s, prec_mode, z, P, W, dlogits, ddist, dsim, dz, dP, dW, StepExtras{}, workspace,
                           workspace_bytes, (cudaStream_t)stream);
}

size_t adlm_head_fwd_big_workspace_bytes(int K, int N, int C) {
  if (K <= 0 || N <= 0 || C <= 0) return 0;
  return tc_head_big_workspace_bytes(K, N, C);
}

int adlm_head_fwd_big(const float* z, int B, int K, int HW, const float* P, const float* W, int N, int C, float eps,
                      int act_mode, float* logits, float* dist, int32_t* argmax, void* workspace,
                      size_t workspace_bytes, void* stream) {
  ADLM_CHECK_ARG(z && P && W && logits, "null pointer argument");
  int rc = check_shape(__func__, B, K, HW, N, C);
  if (rc != ADLM_OK) return rc;
  ADLM_CHECK_ARG(act_mode == ADLM_ACT_LOG || act_mode == ADLM_ACT_LINEAR, "unknown act_mode");
  HeadShape s{B, K, HW, N, C, eps, act_mode};
  return tc_head_fwd_big(s, z, P, W, logits, dist, argmax, workspace, workspace_bytes, (cudaStream_t)stream);
}

size_t adlm_head_bwd_workspace_bytes(int K, int N, int prec_mode) { return head_bwd_workspace_bytes(prec_mode, K, N); }

size_t adlm_step_workspace_bytes(int B, int K, int HW, int N, int C) {
  if (B <= 0 || K <= 0 || HW <= 0 || N <= 0 || C <= 0) return 0;
  return step_workspace_bytes(B, K, HW, N, C);
}

int adlm_step_fwd_bwd(const float* z, int B, int K, int HW, const float* P, const float* W, int N, int C,
                      float eps, int prec_mode, const int32_t* target, const int32_t* proto_class,
                      int ignore_void, float w_ce, float w_kld, float w_l1, float grad_scale,
                      const int32_t* global_counts,
                      float* logits_out, int32_t* argmax_out, float* dz, float* dP, float* dW,
                      double* scalars, void* workspace, size_t workspace_bytes, void* stream) {
  int rc = check_shape(__func__, B, K, HW, N, C);
  if (rc != ADLM_OK) return rc;
  ADLM_CHECK_ARG(z && P && W && target && proto_class && scalars && workspace, "null pointer argument");
  ADLM_CHECK_ARG(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  if (workspace_bytes < step_workspace_bytes(B, K, HW, N, C)) {
    set_error("%s: workspace too small (%zu < %zu)", __func__, workspace_bytes, step_workspace_bytes(B, K, HW, N, C));
    return ADLM_E_WORKSPACE;
  }
  HeadShape s{B, K, HW, N, C, eps, ADLM_ACT_LOG};
  return step_fwd_bwd_impl(s, prec_mode, z, P, W, target, proto_class, ignore_void, w_ce, w_kld, w_l1,
                           grad_scale, global_counts, logits_out, argmax_out, dz, dP, dW, scalars, workspace,
                           nullptr, nullptr, (cudaStream_t)stream);
}

size_t adlm_dp_exchange_bytes(int K, int N, int C, int world) {
  if (K <= 0 || N <= 0 || C <= 0 || world <= 0 || world > ADLM_DP_MAX_WORLD) return 0;
  return dp_layout(K, N, C, world).total;
}

int adlm_dp_exchange_layout(int K, int N, int C, int world, size_t* out8) {
  ADLM_CHECK_ARG(out8 != nullptr, "null pointer argument");
  ADLM_CHECK_ARG(K > 0 && N > 0 && C > 0 && world > 0 && world <= ADLM_DP_MAX_WORLD, "bad size");
  const DpLayout l = dp_layout(K, N, C, world);
  out8[0] = l.cnt; out8[1] = l.cflag; out8[2] = l.acc; out8[3] = l.gflag; out8[4] = l.grad; out8[5] = l.total;
  out8[6] = (size_t)l.n_grad_pad; out8[7] = (size_t)l.n_blocks;
  return ADLM_OK;
}

int adlm_step_fwd_bwd_dp(const float* z, int B, int K, int HW, const float* P, const float* W, int N, int C,
                         float eps, int prec_mode, const int32_t* target, const int32_t* proto_class,
                         int ignore_void, float w_ce, float w_kld, float w_l1, float grad_scale,
                         float* logits_out, int32_t* argmax_out, float* dz, float* dP, float* dW,
                         double* scalars, void* workspace, size_t workspace_bytes, const adlm_dp_comm* comm,
                         void* stream) {
  int rc = check_shape(__func__, B, K, HW, N, C);
  if (rc != ADLM_OK) return rc;
  ADLM_CHECK_ARG(z && P && W && target && proto_class && scalars && workspace && comm, "null pointer argument");
  ADLM_CHECK_ARG(comm->peers != nullptr, "comm->peers is null");
  ADLM_CHECK_ARG(comm->world >= 1 && comm->world <= ADLM_DP_MAX_WORLD, "comm->world outside [1, ADLM_DP_MAX_WORLD]");
  ADLM_CHECK_ARG(comm->rank >= 0 && comm->rank < comm->world, "comm->rank outside [0, world)");
  ADLM_CHECK_ARG(comm->seq >= 1, "comm->seq starts at 1");
  ADLM_CHECK_ARG(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  if (workspace_bytes < step_workspace_bytes(B, K, HW, N, C)) {
    set_error("%s: workspace too small (%zu < %zu)", __func__, workspace_bytes, step_workspace_bytes(B, K, HW, N, C));
    return ADLM_E_WORKSPACE;
  }
  DpComm cm;
  cm.rank = comm->rank;
  cm.world = comm->world;
  cm.timeout_ns = (unsigned long long)(comm->timeout_ms ? comm->timeout_ms : 30000u) * 1000000ull;
  // ADLM_DP_DEFER_COUNTS=1 moves the loss-normaliser rendezvous from k_step_prepare into the forward's first logits
  // epilogue.  Off by default: on 8 GPUs it only moves the wait (prepare -5 us, forward +8 us; 200.8 vs 197.9 us per
  // step on one box, 224 vs 228 us on another) — what the step waits for is the slowest rank, not a latency
  cm.defer_counts = getenv("ADLM_DP_DEFER_COUNTS") != nullptr ? 1 : 0;
  // the flags hold the low 32 bits of the counter; 0 is the buffers' initial state, so it is skipped
  cm.seq = (unsigned int)(comm->seq & 0xffffffffull);
  if (cm.seq == 0u) cm.seq = 0x80000000u | (unsigned int)(comm->seq & 1ull);
  if ((cm.seq & 1u) != (unsigned int)(comm->seq & 1ull)) cm.seq ^= 1u;
  for (int r = 0; r < cm.world; ++r) {
    ADLM_CHECK_ARG(comm->peers[r] != nullptr && ((uintptr_t)comm->peers[r] & 255) == 0,
                   "comm->peers[r] must be non-null and 256-byte aligned");
    cm.peer[r] = static_cast<char*>(comm->peers[r]);
  }
  HeadShape s{B, K, HW, N, C, eps, ADLM_ACT_LOG};
  return step_fwd_bwd_impl(s, prec_mode, z, P, W, target, proto_class, ignore_void, w_ce, w_kld, w_l1,
                           grad_scale, nullptr, logits_out, argmax_out, dz, dP, dW, scalars, workspace,
                           &cm, nullptr, (cudaStream_t)stream);
}

int adlm_head_fwd_pre(const float* pre, int B, int K, int HW, const float* P, const float* W, int N, int C,
                      float eps, int act_mode, int prec_mode, float* z_out, float* logits, float* dist, float* sim,
                      int32_t* argmax, void* stream) {
  int rc = check_shape(__func__, B, K, HW, N, C);
  if (rc != ADLM_OK) return rc;
  ADLM_CHECK_ARG(pre && P && W && logits && z_out, "pre, P, W, z_out and logits must be non-null");
  ADLM_CHECK_ARG(act_mode == ADLM_ACT_LOG || act_mode == ADLM_ACT_LINEAR, "unknown act_mode");
  HeadShape s{B, K, HW, N, C, eps, act_mode};
  if (!head_fuses_sigmoid(s, prec_mode)) {
    set_error("%s: unsupported: needs ADLM_PREC_TF32, K <= 64, N <= 192, C <= 32", __func__);
    return ADLM_E_UNSUPPORTED;
  }
  StepExtras ex;
  ex.z_out = z_out;
  return head_fwd_dispatch(s, prec_mode, pre, P, W, logits, dist, sim, argmax, ex, (cudaStream_t)stream);
}

int adlm_step_fwd_bwd_pre(const float* pre, int B, int K, int HW, const float* P, const float* W, int N, int C,
                          float eps, int prec_mode, const int32_t* target, const int32_t* proto_class,
                          int ignore_void, float w_ce, float w_kld, float w_l1, float grad_scale,
                          const int32_t* global_counts, float* z_out, float* logits_out, int32_t* argmax_out,
                          float* dpre, float* dP, float* dW, double* scalars, void* workspace,
                          size_t workspace_bytes, void* stream) {
  int rc = check_shape(__func__, B, K, HW, N, C);
  if (rc != ADLM_OK) return rc;
  ADLM_CHECK_ARG(pre && P && W && target && proto_class && scalars && workspace && z_out, "null pointer argument");
  ADLM_CHECK_ARG(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  if (workspace_bytes < step_workspace_bytes(B, K, HW, N, C)) {
    set_error("%s: workspace too small (%zu < %zu)", __func__, workspace_bytes, step_workspace_bytes(B, K, HW, N, C));
    return ADLM_E_WORKSPACE;
  }
  HeadShape s{B, K, HW, N, C, eps, ADLM_ACT_LOG};
  return step_fwd_bwd_impl(s, prec_mode, pre, P, W, target, proto_class, ignore_void, w_ce, w_kld, w_l1,
                           grad_scale, global_counts, logits_out, argmax_out, dpre, dP, dW, scalars, workspace,
                           nullptr, z_out, (cudaStream_t)stream);
}

int adlm_step_counts(const int32_t* target, int B, int HW, const int32_t* proto_class, int N, int C,
                     int ignore_void, int32_t* counts_out, void* workspace, size_t workspace_bytes,
                     void* stream) {
  ADLM_CHECK_ARG(target && proto_class && counts_out && workspace, "null pointer argument");
  ADLM_CHECK_ARG(B > 0 && HW > 0 && N > 0 && C > 0, "non-positive size");
  ADLM_CHECK_SUPPORTED(C <= ADLM_MAX_C, "C > 128");
  ADLM_CHECK_ARG(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  if (workspace_bytes < step_workspace_bytes(B, 1, HW, N, C)) {
    set_error("%s: workspace too small", __func__);
    return ADLM_E_WORKSPACE;
  }
  return step_counts_impl(target, B, HW, proto_class, N, C, ignore_void, counts_out, workspace,
                          (cudaStream_t)stream);
}

int adlm_step_counts_noalloc(const int32_t* target, int B, int HW, const int32_t* proto_class, int N, int C,
                             int ignore_void, int32_t* counts_out, void* stream) {
  ADLM_CHECK_ARG(target && proto_class && counts_out, "null pointer argument");
  ADLM_CHECK_ARG(B > 0 && HW > 0 && N > 0 && C > 0, "non-positive size");
  return step_counts_noalloc_impl(target, B, HW, proto_class, N, C, ignore_void, counts_out, (cudaStream_t)stream);
}

int adlm_eval_upsample(const float* logits, const float* dist, int B, int h, int w, int C, int N, int H, int Wimg,
                       const int32_t* y0, const float* ly, const int32_t* x0, const float* lx, int patch_rows,
                       int patch_cols, uint8_t* pred, int16_t* nearest, const uint8_t* labels,
                       const int32_t* proto_class, unsigned long long* counters, void* stream) {
  ADLM_CHECK_ARG(logits && y0 && ly && x0 && lx, "null pointer argument");
  ADLM_CHECK_ARG(B > 0 && h > 0 && w > 0 && C > 0 && H > 0 && Wimg > 0, "non-positive size");
  ADLM_CHECK_ARG(dist == nullptr || N > 0, "N must be positive when distances are given");
  ADLM_CHECK_ARG(patch_rows > 0 && patch_cols > 0 && patch_rows <= h && patch_cols <= w, "bad source patch size");
  ADLM_CHECK_SUPPORTED(C <= 256 && N <= 32767, "C > 256 or N > 32767");
  ADLM_CHECK_SUPPORTED(B <= 65535, "B > 65535 per call");
  if (dist == nullptr) N = 0;
  return eval_upsample_impl(logits, dist, B, h, w, C, N, H, Wimg, y0, ly, x0, lx, patch_rows, patch_cols, pred,
                            nearest, labels, proto_class, counters, (cudaStream_t)stream);
}

int adlm_patch_class_mask(const uint8_t* labels, int n_img, int H, int Wimg, const int32_t* row_lut,
                          const int32_t* col_lut, int h, int w, uint32_t* mask_out, void* stream) {
  ADLM_CHECK_ARG(labels && row_lut && col_lut && mask_out, "null pointer argument");
  ADLM_CHECK_ARG(n_img > 0 && H > 0 && Wimg > 0 && h > 0 && w > 0, "non-positive size");
  ADLM_CHECK_SUPPORTED(n_img <= 65535, "n_img > 65535 per call (split the shard)");
  ADLM_CHECK_SUPPORTED((size_t)w * sizeof(uint32_t) <= 48 * 1024, "feature map wider than 12288 patches");
  return patch_class_mask_impl(labels, n_img, H, Wimg, row_lut, col_lut, h, w, mask_out, (cudaStream_t)stream);
}

int adlm_push_scan(const float* z, int n_img, int K, int HW, const uint32_t* mask, const float* P, int N,
                   const int32_t* proto_class, long long patch_base, int prec_mode,
                   unsigned long long* packed_min, void* stream) {
  ADLM_CHECK_ARG(z && mask && P && proto_class && packed_min, "null pointer argument");
  ADLM_CHECK_ARG(n_img > 0 && K > 0 && HW > 0 && N > 0 && patch_base >= 0, "bad size");
  ADLM_CHECK_SUPPORTED(K <= ADLM_MAX_K, "K > 256");
  ADLM_CHECK_SUPPORTED(patch_base + (long long)n_img * HW < (1ll << 32), "global patch id needs more than 32 bits");
  switch (prec_mode) {
    case ADLM_PREC_FP32:
      return simt_push_scan(z, n_img, K, HW, mask, P, N, proto_class, patch_base, packed_min, 0,
                            (cudaStream_t)stream);
    case ADLM_PREC_TF32:
    case ADLM_PREC_TF32X3:
      return tc_push_scan(z, n_img, K, HW, mask, P, N, proto_class, patch_base, prec_mode, packed_min, 0,
                          (cudaStream_t)stream);
    default:
      set_error("unknown prec_mode %d", prec_mode);
      return ADLM_E_INVALID;
  }
}

int adlm_push_scan_images(const float* z, int n_img, int K, int HW, const uint32_t* mask, const float* P, int N,
                          const int32_t* proto_class, int prec_mode, unsigned long long* packed_min, void* stream) {
  ADLM_CHECK_ARG(z && P && packed_min, "null pointer argument");
  ADLM_CHECK_ARG(n_img > 0 && K > 0 && HW > 0 && N > 0, "bad size");
  ADLM_CHECK_SUPPORTED(K <= ADLM_MAX_K, "K > 256");
  ADLM_CHECK_SUPPORTED((long long)n_img * HW < (1ll << 31), "more than 2^31 patches per call (split the batch)");
  switch (prec_mode) {
    case ADLM_PREC_FP32:
      return simt_push_scan(z, n_img, K, HW, mask, P, N, proto_class, 0, packed_min, 1, (cudaStream_t)stream);
    case ADLM_PREC_TF32:
    case ADLM_PREC_TF32X3:
      return tc_push_scan(z, n_img, K, HW, mask, P, N, proto_class, 0, prec_mode, packed_min, 1,
                          (cudaStream_t)stream);
    default:
      set_error("unknown prec_mode %d", prec_mode);
      return ADLM_E_INVALID;
  }
}

int adlm_min_pool_fwd(const float* z, int B, int K, int HW, const float* P, int N, int prec_mode,
                      unsigned long long* packed, float* min_dist, int32_t* argmin, void* stream) {
  ADLM_CHECK_ARG(min_dist != nullptr, "null pointer argument");
  ADLM_CHECK_SUPPORTED((long long)B * N < (1ll << 31), "B*N too large");
  const int rc = adlm_push_scan_images(z, B, K, HW, nullptr, P, N, nullptr, prec_mode, packed, stream);
  if (rc != ADLM_OK) return rc;
  return min_pool_unpack_impl(packed, B, N, min_dist, argmin, (cudaStream_t)stream);
}

int adlm_min_pool_bwd(const float* z, int B, int K, int HW, const float* P, int N,
                      const unsigned long long* packed, const float* grad_min, float* dz, float* dP, void* stream) {
  ADLM_CHECK_ARG(z && P && packed && grad_min, "null pointer argument");
  ADLM_CHECK_ARG(B > 0 && K > 0 && HW > 0 && N > 0, "bad size");
  ADLM_CHECK_SUPPORTED(B <= 65535, "B > 65535 per call");
  if (!dz && !dP) return ADLM_OK;
  return min_pool_bwd_impl(z, B, K, HW, P, N, packed, grad_min, dz, dP, (cudaStream_t)stream);
}

int adlm_push_gather(const float* z, int n_img, int K, int HW, long long patch_base,
                     const unsigned long long* packed_min, int N, float* proto_out, void* stream) {
  ADLM_CHECK_ARG(z && packed_min && proto_out, "null pointer argument");
  ADLM_CHECK_ARG(n_img > 0 && K > 0 && HW > 0 && N > 0 && patch_base >= 0, "bad size");
  return push_gather_impl(z, n_img, K, HW, patch_base, packed_min, N, proto_out, (cudaStream_t)stream);
}

}  // extern "C"
