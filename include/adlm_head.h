/*
 * adlm_head.h — C ABI of the B200-native ProtoSeg prototype head + push search.
 *
 * One shared library (libadlm_head.so, built from
 * adlm_inherently_interpretable_segmentation_b200/csrc/ by nvcc for sm_100a) exports exactly
 * the symbols declared here.  Plain C: pointers + sizes, no torch types.  Every pointer is a
 * DEVICE pointer unless it says "host"; every buffer is allocated and owned by the caller
 * (torch, in the Python host); nothing is allocated inside.  Every call is asynchronous on `stream` (a cudaStream_t passed as void*).
 * Return value: 0 on success, negative ADLM_E_* on failure, message via adlm_last_error().
 * There is no CPU fallback and no fallback to a library GEMM: unsupported shapes return
 * ADLM_E_UNSUPPORTED.
 *
 * Reference interface replaced (file:line into the reference repo):
 *   adlm_head_fwd          model.py:203-221 (_l2_convolution), :231-237 (distance_2_similarity),
 *                          :161-162 (run_last_layer), :259-283 (forward_from_conv_features)
 *   adlm_head_bwd          autograd backward of the above, invoked at segmentation/module.py:236
 *   adlm_step_fwd_bwd      segmentation/module.py:137-222 + :236 (forward, CE on non-void patches,
 *                          KLD prototype-diversity loss on distances, masked L1, backward)
 *   adlm_step_fwd_bwd_dp   the same step with the batch sharded over GPUs (the reference trains with
 *                          Trainer.gpus = 1, segmentation/configs/cityscapes_kld_imnet.gin:1; raising it makes
 *                          Lightning wrap the module in DDP, whose gradient all-reduce this entry point
 *                          performs inside the step's own kernels over peer-mapped memory)
 *   adlm_head_fwd_pre, adlm_step_fwd_bwd_pre   the same with model.py:124-128,174 (Sigmoid add-on layer) fused in
 *   adlm_eval_upsample     segmentation/eval_valid.py:171-199, eval_test.py:105-106, push.py:243-247 (bilinear
 *                          upsampling of logits / distances to image size + argmax / argmin + IoU counters)
 *   adlm_patch_class_mask  segmentation/push.py:201-230 (pixel -> patch class sets)
 *   adlm_push_scan         segmentation/push.py:249-280 (class-filtered per-prototype argmin,
 *                          strict '<' against the global minimum)
 *   adlm_push_gather       segmentation/push.py:268-280, :142-145 (winner feature vectors)
 *   adlm_push_scan_images  find_nearest.py:128-160 (per-image, per-prototype nearest patch; void-masked)
 *   adlm_min_pool_fwd/bwd  model.py:284-299 (legacy classification branch: global min-pool of the distance
 *                          map) and its autograd backward, invoked at train_and_test.py:101
 */
#ifndef ADLM_HEAD_H
#define ADLM_HEAD_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ADLM_ABI_VERSION 9

#if defined(__GNUC__)
#define ADLM_API __attribute__((visibility("default")))
#else
#define ADLM_API
#endif

/* error codes */
#define ADLM_OK 0
#define ADLM_E_INVALID (-1)     /* bad argument (null pointer, non-positive size, misaligned) */
#define ADLM_E_UNSUPPORTED (-2) /* shape / mode outside what the kernels implement */
#define ADLM_E_CUDA (-3)        /* a CUDA runtime call failed; message holds cudaGetErrorString */
#define ADLM_E_WORKSPACE (-4)   /* workspace too small */

/* prototype_activation_function (model.py:231-237) */
#define ADLM_ACT_LOG 0    /* s = log((d + 1) / (d + eps)) */
#define ADLM_ACT_LINEAR 1 /* s = -d */

/* arithmetic of the z.p contraction */
#define ADLM_PREC_FP32 0   /* CUDA-core fp32 FMA — the reference's arithmetic class */
#define ADLM_PREC_TF32 1   /* tcgen05 kind::tf32, operands centred and rounded to nearest */
#define ADLM_PREC_TF32X3 2 /* tcgen05 3xTF32 (hi*hi + hi*lo + lo*hi): fp32-class accuracy */

/* limits of this build (checked at run time, ADLM_E_UNSUPPORTED beyond them) */
#define ADLM_MAX_K 256            /* prototype depth */
#define ADLM_MAX_C 128            /* classes */
#define ADLM_MAX_PROTOS_PER_CLASS 16
#define ADLM_MAX_PUSH_CLASSES 32  /* patch class mask is one uint32 per patch */
#define ADLM_DP_MAX_WORLD 16       /* ranks of one NVLink domain in adlm_step_fwd_bwd_dp */

/* number of doubles adlm_step_fwd_bwd writes to `scalars` */
#define ADLM_STEP_NSCALARS 8
/* scalars[0] loss  [1] cross_entropy  [2] kld_loss  [3] l1  [4] n_correct  [5] n_patches (non-void)
 * [6] n_kld_terms  [7] 0, or > 0 when the step failed: a peer exchange of adlm_step_fwd_bwd_dp timed out (results are
 * NaN), or a class has more than ADLM_MAX_PROTOS_PER_CLASS prototypes (the KLD term is then incomplete) */

ADLM_API const char* adlm_last_error(void);
ADLM_API int adlm_abi_version(void);
/* number of kernel launches issued by this library since load (bench.py's gpu_launches) */
ADLM_API unsigned long long adlm_launch_count(void);

/*
 * Measurement aid (bench.py's roofline leg): between adlm_timing_begin(tag, max) and adlm_timing_end the
 * library brackets every launch of the tagged kernel with CUDA events on the launching stream.
 * adlm_timing_end synchronises those events and writes the per-launch durations (milliseconds) to the
 * HOST array ms_out (capacity given to begin); returns the number of launches recorded or a negative code.
 */
#define ADLM_K_HEAD_FWD 1
#define ADLM_K_HEAD_BWD 2
#define ADLM_K_PUSH_SCAN 3
#define ADLM_K_PATCH_MASK 4
#define ADLM_K_KLD 5
#define ADLM_K_STEP_PREPARE 6
ADLM_API int adlm_timing_begin(int kernel_tag, int max_samples);
ADLM_API int adlm_timing_end(float* ms_out);

/*
 * Head forward.  z [B,K,HW] fp32 (NCHW features flattened over h*w, contiguous),
 * P [N,K] prototypes, W [C,N] last layer (no bias).
 * Outputs (each may be NULL = not wanted, except logits):
 *   logits [B,HW,C]   (= reference (B,h,w,C))
 *   dist   [B,N,HW]   (= reference (B,N,h,w)); the M x N map touches HBM only if requested
 *   sim    [B*HW,N]   (= reference prototype_activations)
 *   argmax [B,HW]     int32 per-pixel argmax over classes (first maximum, as torch.argmax)
 */
ADLM_API int adlm_head_fwd(const float* z, int B, int K, int HW,
                  const float* P, const float* W, int N, int C,
                  float eps, int act_mode, int prec_mode,
                  float* logits, float* dist, float* sim, int32_t* argmax,
                  void* stream);

/*
 * Head forward for LARGE heads on the tensor cores (model.py:203-283 for any prototype_shape, model.py:54: BASELINE
 * configs[4], 1000 prototypes x 256-d, 100 classes): K <= 256, N <= 1024, C <= 112, single-pass TF32 on centred
 * operands like ADLM_PREC_TF32 of adlm_head_fwd.  The kernel runs on CTA pairs (clusters of two, tcgen05 cta_group::2):
 * each CTA keeps one 128-pixel tile, the pair shares every prototype / last-layer chunk, which streams through shared
 * memory by TMA bulk copies from a copy of P / W that a small kernel of the same call stages in `workspace`
 * (adlm_head_fwd_big_workspace_bytes(K, N, C) bytes, 256-byte aligned; contents need not survive the call).
 * logits [B,HW,C]; dist [B,N,HW] | NULL (at configs[4]'s full size the map is 8.4 GB per image: ask for it on crops
 * only); argmax [B,HW] | NULL.  Larger shapes: ADLM_E_UNSUPPORTED (use adlm_head_fwd with ADLM_PREC_FP32).
 */
ADLM_API size_t adlm_head_fwd_big_workspace_bytes(int K, int N, int C);
ADLM_API int adlm_head_fwd_big(const float* z, int B, int K, int HW,
                      const float* P, const float* W, int N, int C,
                      float eps, int act_mode,
                      float* logits, float* dist, int32_t* argmax,
                      void* workspace, size_t workspace_bytes, void* stream);

/*
 * Head backward for arbitrary upstream gradients (autograd drop-in).
 *   dlogits [B,HW,C] | NULL, ddist [B,N,HW] | NULL, dsim [B*HW,N] | NULL
 *   dz [B,K,HW] | NULL, dP [N,K] | NULL, dW [C,N] | NULL  (dP/dW are overwritten, not accumulated)
 * Distances are recomputed from z and P (never read back from HBM).
 * ADLM_PREC_TF32 with K <= 64, N <= 192, C <= 32 runs on the tensor cores (two kernels: dZ with pixels in the
 * TMEM lanes, dP/dW with prototypes in the lanes); 3xTF32 and larger shapes use the fp32 CUDA-core backward.
 */
ADLM_API int adlm_head_bwd(const float* z, int B, int K, int HW,
                  const float* P, const float* W, int N, int C,
                  float eps, int act_mode, int prec_mode,
                  const float* dlogits, const float* ddist, const float* dsim,
                  float* dz, float* dP, float* dW,
                  void* workspace, size_t workspace_bytes,
                  void* stream);
/* scratch adlm_head_bwd needs for (K, N, prec_mode): per-CTA partial sums of dP/dW of the tensor-core
 * backward (0 for the modes / shapes that use the fp32 backward; workspace may then be NULL) */
ADLM_API size_t adlm_head_bwd_workspace_bytes(int K, int N, int prec_mode);

/*
 * Fused training / validation step of the head ('log' activation):
 *   forward -> CE over non-void patches -> KLD over same-class prototype pairs (on distances,
 *   groups keyed by the UN-shifted label, module.py:150,170-178) -> masked L1 on W -> backward.
 *   target [B,HW] int32: labels already resized to the feature map (dataset.py:23-31);
 *     ignore_void != 0: label 0 is void, class = label-1 (module.py:156-160)
 *   proto_class [N] int32: class of each prototype (argmax of prototype_class_identity row)
 *   grad_scale: multiplies every gradient (1/iter_size, module.py:236)
 *   dz/dP/dW: NULL, NULL, NULL = loss/metrics only (validation step).  dP/dW overwritten.
 *   logits_out [B,HW,C] | NULL, argmax_out [B,HW] | NULL.
 *   scalars: ADLM_STEP_NSCALARS doubles (device).
 *   workspace: adlm_step_workspace_bytes(...) bytes (device, 256-byte aligned).
 */
ADLM_API size_t adlm_step_workspace_bytes(int B, int K, int HW, int N, int C);
ADLM_API int adlm_step_fwd_bwd(const float* z, int B, int K, int HW,
                      const float* P, const float* W, int N, int C,
                      float eps, int prec_mode,
                      const int32_t* target, const int32_t* proto_class, int ignore_void,
                      float w_ce, float w_kld, float w_l1, float grad_scale,
                      const int32_t* global_counts,
                      float* logits_out, int32_t* argmax_out,
                      float* dz, float* dP, float* dW,
                      double* scalars, void* workspace, size_t workspace_bytes,
                      void* stream);

/*
 * The same two entry points with the add-on activation fused in (model.py:124-128,174: add_on_layers =
 * Sequential(Sigmoid())).  `pre` [B,K,HW] holds the PRE-sigmoid features; the forward
 * applies the sigmoid while it stages a tile and writes z = sigmoid(pre) to z_out [B,K,HW] (the reference hands z to
 * forward_with_features / push, and the backward kernels re-read it); `dpre` receives the gradient wrt `pre`
 * (dz * z * (1 - z), applied in the dZ kernel's store).  Saves the two elementwise passes around the head.
 * Tensor-core path only: ADLM_PREC_TF32, K <= 64, N <= 192, C <= 32, else ADLM_E_UNSUPPORTED (apply the sigmoid
 * with stock code and call the plain entry points).
 */
ADLM_API int adlm_head_fwd_pre(const float* pre, int B, int K, int HW,
                      const float* P, const float* W, int N, int C,
                      float eps, int act_mode, int prec_mode,
                      float* z_out, float* logits, float* dist, float* sim, int32_t* argmax,
                      void* stream);
ADLM_API int adlm_step_fwd_bwd_pre(const float* pre, int B, int K, int HW,
                          const float* P, const float* W, int N, int C,
                          float eps, int prec_mode,
                          const int32_t* target, const int32_t* proto_class, int ignore_void,
                          float w_ce, float w_kld, float w_l1, float grad_scale,
                          const int32_t* global_counts,
                          float* z_out, float* logits_out, int32_t* argmax_out,
                          float* dpre, float* dP, float* dW,
                          double* scalars, void* workspace, size_t workspace_bytes,
                          void* stream);

/*
 * Loss normalisers of a batch shard: counts_out[0] = number of CE patches (module.py:156-165),
 * counts_out[1] = number of KLD terms (module.py:186-206).  Data-parallel training sums these over
 * ranks and passes the sums back as `global_counts` (device, 2 x int32; NULL = use the shard's own),
 * so that the summed shard gradients equal the single-GPU gradient of the whole batch.
 */
ADLM_API int adlm_step_counts(const int32_t* target, int B, int HW,
                     const int32_t* proto_class, int N, int C, int ignore_void,
                     int32_t* counts_out, void* workspace, size_t workspace_bytes, void* stream);

/*
 * adlm_step_counts without a workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same inputs and the same counts_out as adlm_step_counts; the kernel's
 * scratch (label histograms, class sizes, prototype ranks) lives in a static device array of
 * ADLM_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.  Limits: B <= ADLM_NOALLOC_MAX_B,
 * N <= ADLM_NOALLOC_MAX_N, C <= ADLM_MAX_C, device index < 64; larger sizes return ADLM_E_UNSUPPORTED (use
 * adlm_step_counts).  All calls on a device share the array, which keeps the last call's contents; nothing in it has
 * to start at zero (the kernel overwrites what it reads), but calls that OVERLAP on the GPU (different streams,
 * or the same stream from two threads without ordering) race on it — serialise them.  The first call on a device
 * looks up the array's device address (cudaGetSymbolAddress); no call allocates or frees memory.
 */
#define ADLM_NOALLOC_MAX_B 256
#define ADLM_NOALLOC_MAX_N 1024
#define ADLM_NOALLOC_SCRATCH_BYTES 136960
ADLM_API int adlm_step_counts_noalloc(const int32_t* target, int B, int HW,
                             const int32_t* proto_class, int N, int C, int ignore_void,
                             int32_t* counts_out, void* stream);

/*
 * Data-parallel fused step over peer-mapped memory (one process per GPU, all GPUs of one NVLink domain).
 * Same arguments and results as adlm_step_fwd_bwd on this rank's image shard, except that
 *   - the two loss normalisers are summed over ranks inside the step's prologue kernel,
 *   - dP, dW and the loss sums are summed over ranks inside the dP/dW partials-reduction kernel:
 *     every rank stores its partial result into its slot of every peer's exchange buffer (NVLink stores),
 *     release-stores a sequence flag, waits for its peers' flags and adds the slots in rank order,
 * so dP, dW and scalars come out bit-identical on all ranks and equal to the single-GPU step on the whole
 * batch; dz is this shard's.  No NCCL call, no extra launch.
 *   comm->peers: HOST array of `world` DEVICE pointers; peers[r] is rank r's exchange buffer as mapped into
 *     this process (CUDA IPC / VMM / torch symmetric memory), each adlm_dp_exchange_bytes(K,N,C,world) bytes,
 *     256-byte aligned, zero-filled once before the first step and never touched by the caller afterwards;
 *   comm->seq: step counter, identical on all ranks, 1 for the first step and +1 per step on this buffer.
 * Requires ADLM_PREC_TF32, K <= 64, N <= 192, C <= 32 and non-NULL dz/dP/dW (else ADLM_E_UNSUPPORTED: use
 * adlm_step_counts + adlm_step_fwd_bwd(global_counts) + a library all-reduce).  A peer that never arrives
 * makes the waits time out after comm->timeout_ms (wall clock): the outputs are then NaN and scalars[7] > 0.
 */
typedef struct adlm_dp_comm {
  int rank;
  int world;
  unsigned long long seq;
  void* const* peers;
  unsigned int timeout_ms; /* bound of every in-kernel wait for a peer; 0 = 30000 */
} adlm_dp_comm;
ADLM_API size_t adlm_dp_exchange_bytes(int K, int N, int C, int world);
/* diagnostics / tests: byte offsets inside one exchange buffer, HOST array of 8:
 * [0] normalisers int32[2][world][2]  [1] their flags u32[2][world]  [2] loss sums f64[2][world][4]
 * [3] gradient flags u32[2][world][n_blocks]  [4] gradients f32[2][world][n_grad_pad]  [5] total bytes
 * [6] n_grad_pad  [7] n_blocks.  First index = seq & 1, second = publishing rank. */
ADLM_API int adlm_dp_exchange_layout(int K, int N, int C, int world, size_t* out8);
ADLM_API int adlm_step_fwd_bwd_dp(const float* z, int B, int K, int HW,
                         const float* P, const float* W, int N, int C,
                         float eps, int prec_mode,
                         const int32_t* target, const int32_t* proto_class, int ignore_void,
                         float w_ce, float w_kld, float w_l1, float grad_scale,
                         float* logits_out, int32_t* argmax_out,
                         float* dz, float* dP, float* dW,
                         double* scalars, void* workspace, size_t workspace_bytes,
                         const adlm_dp_comm* comm, void* stream);

/*
 * Evaluation epilogue: F.interpolate(mode='bilinear', align_corners=False) of the
 * logits and of the prototype distances to the image size, fused with torch.argmax / torch.argmin over the planes
 * and with the reference's counters; the upsampled stacks are never materialised.
 *   logits [B,h,w,C] (adlm_head_fwd layout), dist [B,N,h,w] | NULL (then nearest / prototype counts are skipped);
 *   y0 [H] int32, ly [H] fp32, x0 [Wimg] int32, lx [Wimg] fp32: source index and lambda of every output row /
 *     column, computed by the caller in FLOAT32 as torch does: src = max(scale * (dst + 0.5) - 0.5, 0) with
 *     scale = (float)in / out, index = (int)src, lambda = src - index  (device arrays);
 *   patch_rows / patch_cols: the largest number of source rows / columns any ADLM_EVAL_TILE_Y x ADLM_EVAL_TILE_X
 *     output tile touches: max over tiles of min(y0[last row of tile] + 1, h - 1) - y0[first row] + 1 (same for x);
 *   pred [B,H,Wimg] uint8 | NULL: argmax class; nearest [B,H,Wimg] int16 | NULL: argmin prototype;
 *   labels [B,H,Wimg] uint8 | NULL: ground truth, 0 = void, class + 1 otherwise;
 *   proto_class [N] int32 | NULL; counters | NULL: 2 + 2C + N unsigned 64-bit, ACCUMULATED (caller zeroes):
 *     [0] correct pixels, [1] non-void pixels, [2..2+C) per-class intersection, [2+C..2+2C) per-class union
 *     (eval_valid.py:178-189), [2+2C..) pixels whose nearest prototype n belongs to the predicted class (:192-197).
 */
#define ADLM_EVAL_TILE_X 64
#define ADLM_EVAL_TILE_Y 16
ADLM_API int adlm_eval_upsample(const float* logits, const float* dist, int B, int h, int w, int C, int N,
                       int H, int Wimg,
                       const int32_t* y0, const float* ly, const int32_t* x0, const float* lx,
                       int patch_rows, int patch_cols,
                       uint8_t* pred, int16_t* nearest, const uint8_t* labels,
                       const int32_t* proto_class, unsigned long long* counters, void* stream);

/*
 * Push, step 1: per-patch class bitmask from full-resolution labels.
 *   labels [n_img,H,Wimg] uint8 (0 = void, class c = label-1);
 *   row_lut [H], col_lut [Wimg] int32: pixel -> patch index, computed on the host with the
 *   reference's float expression int(pixel / (H / h)) (push.py:201-202,217,219);
 *   mask_out [n_img,h*w] uint32, bit c set iff >= 1 pixel of class c falls in the patch.
 */
ADLM_API int adlm_patch_class_mask(const uint8_t* labels, int n_img, int H, int Wimg,
                          const int32_t* row_lut, const int32_t* col_lut, int h, int w,
                          uint32_t* mask_out, void* stream);

/*
 * Push, step 2: streaming class-filtered nearest-patch search.
 *   z [n_img,K,HW], mask [n_img,HW], P [N,K], proto_class [N].
 *   packed_min [N] uint64, updated in place with
 *       min(packed_min[j], (float_bits(d) << 32) | (patch_base + img*HW + patch))
 *   over patches whose mask has bit proto_class[j].  d >= 0, so the IEEE bits order like the
 *   values; ties resolve to the lowest global patch id (= earliest image, push.py:265).
 *   Initialise with ADLM_PUSH_NONE.  patch_base + n_img*HW must be < 2^32.
 */
#define ADLM_PUSH_NONE 0x7fffffffffffffffULL
ADLM_API int adlm_push_scan(const float* z, int n_img, int K, int HW,
                   const uint32_t* mask, const float* P, int N, const int32_t* proto_class,
                   long long patch_base, int prec_mode,
                   unsigned long long* packed_min, void* stream);

/*
 * Per-image variant of the scan (find_nearest.py:128-160 takes the per-prototype minimum of every image's
 * distance map; model.py:284-288 min-pools it): one launch for the whole batch.
 *   packed_min [n_img,N] uint64, updated in place with
 *       min(packed_min[i][j], (float_bits(d) << 32) | patch)       patch = index inside image i
 *   mask may be NULL (every patch is a candidate), proto_class may be NULL (no class filter: every
 *   prototype reads mask bit 0).  Initialise with ADLM_PUSH_NONE.
 */
ADLM_API int adlm_push_scan_images(const float* z, int n_img, int K, int HW,
                          const uint32_t* mask, const float* P, int N, const int32_t* proto_class,
                          int prec_mode, unsigned long long* packed_min, void* stream);

/*
 * Legacy classification branch (patch_classification = False, model.py:284-299): global min-pool of the
 * distance map without materialising it.
 *   fwd: packed [B,N] uint64 workspace (initialised with ADLM_PUSH_NONE by the caller, kept for the
 *        backward), min_dist [B,N] = min over h*w of relu(|z|^2 - 2 z.p + |p|^2), argmin [B,N] int32
 *        (patch index, may be NULL).
 *   bwd: grad_min [B,N] -> dz [B,K,HW] and dP [N,K], both ACCUMULATED with atomics (caller zero-fills;
 *        either may be NULL): the gradient reaches only the arg-min patch (max_pool2d backward,
 *        model.py:287) and only where d > 0 (relu, model.py:218).
 */
ADLM_API int adlm_min_pool_fwd(const float* z, int B, int K, int HW, const float* P, int N, int prec_mode,
                      unsigned long long* packed, float* min_dist, int32_t* argmin, void* stream);
ADLM_API int adlm_min_pool_bwd(const float* z, int B, int K, int HW, const float* P, int N,
                      const unsigned long long* packed, const float* grad_min,
                      float* dz, float* dP, void* stream);

/*
 * Push, step 3: fetch the winners that live in this shard.
 *   For each prototype j whose packed id falls in [patch_base, patch_base + n_img*HW):
 *   proto_out[j,:] = z[img,:,patch]; otherwise the row is left untouched (caller zero-fills:
 *   prototypes with no candidate keep the zero vector, push.py:57-61).
 */
ADLM_API int adlm_push_gather(const float* z, int n_img, int K, int HW, long long patch_base,
                     const unsigned long long* packed_min, int N,
                     float* proto_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ADLM_HEAD_H */
