// tcgen05 (5th-gen tensor core) kernels of the prototype head: ADLM_PREC_TF32 / ADLM_PREC_TF32X3.
#pragma once
#include "common.cuh"

namespace adlm {

struct TcFwdParams {
  const float* z;
  const float* P;
  const float* W;
  float* logits;
  float* dist;
  float* sim;
  int32_t* argmax;
  const int32_t* target;
  const int32_t* proto_class;
  const int32_t* proto_rank;
  float* own_dist;
  // fused cross entropy (warp-specialised kernel only): all null / 0 = plain forward
  const int32_t* ce_counts;   // [0] = number of CE patches (device)
  DpComm ce_dp;               // seq != 0: collect the peers' counts here (StepExtras::ce_dp)
  DpLayout ce_dp_lay;
  int32_t* ce_counts_out;     // [2] global sums for the kernels after the forward (written by CTA 0)
  float* dlogits;             // [M, C] | null
  double* ce_accum;           // [0] += sum of losses, [1] += correct patches
  float ce_scale;             // w_ce * grad_scale
  int ignore_void;
  float* z_out;               // [B,K,HW] | null: input is pre-sigmoid, z = sigmoid(input) is written here
  int B, K, HW, N, C;
  float eps;
  int act_mode;
  int Npad, Cpad;
  long long M;
  int num_tiles;
  // launched right after k_step_prepare with programmatic stream serialisation: the CTAs start while the prologue
  // kernel still runs, stage their operands and only then wait for it (griddepcontrol.wait)
  int after_prepare;
  // 3xTF32 forward (k_tc_fwd<.., true>): last layer on the tensor cores too (S and W split in hi + lo, three MMAs per
  // k-step) when the doubled last-layer operand fits shared memory; otherwise fp32 FMAs in the epilogue threads
  int x3_tensor_last;
};
// 1xTF32, K <= 64: producer / MMA / epilogue warps, double-buffered TMEM accumulators (tc_head_ws.cu)
int tc_head_fwd_ws(const TcFwdParams& prm, cudaStream_t st);

int tc_head_fwd(const HeadShape& s, int prec_mode, const float* z, const float* P, const float* W,
                float* logits, float* dist, float* sim, int32_t* argmax, const StepExtras& ex,
                cudaStream_t st);
int tc_head_bwd(const HeadShape& s, int prec_mode, const float* z, const float* P, const float* W,
                const float* dlogits, const float* ddist, const float* dsim, float* dz, float* dP,
                float* dW, const StepExtras& ex, void* workspace, size_t workspace_bytes, cudaStream_t st);
// bytes of scratch tc_head_bwd needs for the per-CTA dP/dW partial sums (0 when the fp32 backward is used)
size_t tc_bwd_workspace_bytes(int K, int N);
int tc_push_scan(const float* z, int n_img, int K, int HW, const uint32_t* mask, const float* P, int N,
                 const int32_t* proto_class, long long patch_base, int prec_mode,
                 unsigned long long* packed_min, int per_image, cudaStream_t st);

}  // namespace adlm
