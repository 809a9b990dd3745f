/*
 * mrisr.h -- C ABI of the B200-native U-Net super-resolution hot path.
 *
 * The reference (rdd0582/mri_superresolution) is pure Python/PyTorch and has no FFI or plugin
 * interface of its own: its seam for this path is the nn.Module API
 *   models/unet_model.py:129,189   UNetSuperRes.__init__/forward
 *   utils/losses.py:27             ssim()
 *   utils/losses.py:159,200        CombinedLoss.__init__/forward
 *   utils/losses.py:242            SSIM
 *   scripts/train.py:186,296-323   Adam + train-step body
 *   scripts/infer.py:268-276       inference body
 * Each entry point below states which of those it replaces.  The Python mirror of the module
 * API (mri_superresolution_b200/models/unet_model.py, utils/losses.py) calls ONLY these
 * functions (ctypes, raw device pointers); INTEGRATION.md shows the binding.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless named host_*; nothing is allocated or freed by
 *     the library: the caller owns all memory, including the workspace.
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), never
 *     synchronises, is re-entrant and CUDA-graph capturable.
 *   - return value: 0 ok; > 0 a cudaError_t; < 0 an MRISR_E_* argument error.
 *     mrisr_last_error() returns a thread-local message for the last non-zero return.
 *   - images are N x 1 x H x W fp32, dense NCHW (== NHWC for one channel).
 *   - parameters / gradients / Adam moments are flat fp32 arenas laid out in the reference's
 *     parameter registration order (models/unet_model.py:129-175); see
 *     mrisr_unet_param_offset().  The checkpoint layout (state_dict keys) is owned by Python.
 */
#ifndef MRISR_H
#define MRISR_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MRISR_VERSION 100

#define MRISR_E_BADARG   (-1)   /* null pointer / non-positive size */
#define MRISR_E_SHAPE    (-2)   /* H or W not a multiple of 8, base_filters not a multiple of 16 */
#define MRISR_E_DTYPE    (-3)   /* unknown precision mode */
#define MRISR_E_WORKSPACE (-4)  /* workspace too small */
#define MRISR_E_STATE    (-5)   /* backward without a matching train-mode forward */

#define MRISR_FP32 0            /* activations fp32, CUDA-core convolutions (parity mode, 1e-4) */
#define MRISR_BF16 1            /* activations/operands bf16, fp32 accumulate, tcgen05 convolutions */

int mrisr_version(void);
const char* mrisr_last_error(void);

/* ---- parameter arena layout (models/unet_model.py:129-175 registration order) ------------- */
/* number of parameter tensors (64) and total fp32 elements (1 823 122 at f = 32) */
int      mrisr_unet_param_tensors(int base_filters);
int64_t  mrisr_unet_param_elems(int base_filters);
/* element offset of tensor `index` in the arena, and its element count */
int64_t  mrisr_unet_param_offset(int base_filters, int index);
int64_t  mrisr_unet_param_size(int base_filters, int index);

/* ---- U-Net forward / backward (replaces UNetSuperRes.forward + its autograd graph) -------- */
/* bytes of caller-owned workspace for one (N,H,W,f,precision) problem.  `train` != 0 keeps every
 * tensor the backward pass needs.  The workspace must be ZERO-FILLED once before its first
 * use with a given shape (zero borders of the padded NHWC activations are an invariant that the
 * kernels preserve) and may then be reused for every later call of the same shape. */
size_t mrisr_unet_workspace_bytes(int N, int H, int W, int base_filters, int precision, int train);

/* x: N x 1 x H x W fp32  ->  y: N x 1 x 2H x 2W fp32 (sigmoid output, models/unet_model.py:211) */
int mrisr_unet_forward(const float* params, const float* x, float* y, void* workspace, size_t workspace_bytes,
                       int N, int H, int W, int base_filters, int precision, int train, void* stream);

/* dy: gradient wrt y (N x 1 x 2H x 2W fp32); y: the forward output itself (sigmoid backward needs it);
 * grads: flat fp32 arena, OVERWRITTEN (not accumulated) with d(loss)/d(param); dx: must be NULL (the
 * reference never asks for the gradient wrt the input image, scripts/train.py:296-318).  Must follow
 * a train-mode mrisr_unet_forward on the same workspace with the same params. */
int mrisr_unet_backward(const float* params, const float* dy, const float* y, float* grads, float* dx,
                        void* workspace, size_t workspace_bytes, int N, int H, int W, int base_filters,
                        int precision, void* stream);

/* ---- data-parallel support: gradient-readiness marks ------------------------------------------------------------
 * The reference is single-device (scripts/train.py:296-318 has no exchange step); BASELINE.json config 3 adds a bucketed
 * gradient all-reduce that overlaps the backward pass.  The backward schedule finishes the conv blocks in the order
 * 19, 17, 18, 16, 15, ..., 0 and parameter offsets grow with the block index, so after block l every arena element at
 * offset >= mrisr_unet_grads_final_from(f, l) is final.  mrisr_unet_backward_marked is mrisr_unet_backward that also
 * records host_mark_events[k] (cudaEvent_t, e.g. from mrisr_event_create) on `stream` as soon as every gradient at
 * offset >= host_mark_offsets[k] is final: a side stream that waits on event k may all-reduce grads[offsets[k] : ) while
 * the rest of the backward pass still runs.  Both arrays are HOST arrays of n_marks (<= 32) entries. */
int64_t mrisr_unet_grads_final_from(int base_filters, int conv_index);
int mrisr_unet_backward_marked(const float* params, const float* dy, const float* y, float* grads, float* dx,
                               void* workspace, size_t workspace_bytes, int N, int H, int W, int base_filters,
                               int precision, void* stream, int n_marks, const int64_t* host_mark_offsets,
                               void* const* host_mark_events);
/* timing-disabled CUDA events owned by the library (NULL on failure), and cudaStreamWaitEvent on a caller stream */
void* mrisr_event_create(void);
int   mrisr_event_destroy(void* event);
int   mrisr_stream_wait_event(void* stream, void* event);

/* ---- fused L1 + SSIM loss and gradient (replaces utils/losses.py:27-81,200-226) ------------ */
/* out5 (device, 5 floats): [0] mean|p-t|  [1] unclamped SSIM  [2] total = l1_w*L1 + ssim_w*(1-clamp(ssim,0,1))
 *                          [3], [4] reserved (perceptual term, total incl. perceptual).
 * dpred (nullable): d(total)/d(pred), N x 1 x H x W fp32.  `scratch` >= mrisr_l1_ssim_scratch_bytes(). */
size_t mrisr_l1_ssim_scratch_bytes(int N, int H, int W);
int mrisr_l1_ssim_fwd_bwd(const float* pred, const float* target, float* dpred, float* out5, void* scratch,
                          int N, int H, int W, float l1_weight, float ssim_weight,
                          int window_size, float sigma, float val_range, void* stream);
/* SSIM metric only (utils/losses.py:242-266): per_sample[n] = mean SSIM map of sample n */
int mrisr_ssim_forward(const float* img1, const float* img2, float* per_sample, void* scratch,
                       int N, int H, int W, int window_size, float sigma, float val_range, void* stream);

/* Allocation-free variant of mrisr_ssim_forward.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Same result as mrisr_ssim_forward; the kernel's block partials live in a file-scope `static __device__` array of
 * mrisr_ssim_static_scratch_bytes() = 1 MiB that every GPU loading the library holds until the process exits, so the call
 * takes no scratch pointer and performs no cudaMalloc / cudaFree / allocating library call.  Limit: 8 bytes per 32 x 32 tile,
 * N * ceil(H/32) * ceil(W/32) * 8 <= 1 MiB (e.g. 128 slices of 1024 x 1024); larger problems return MRISR_E_SHAPE with a
 * message naming the limit.  All calls on a device share the array: calls that overlap on the GPU (different streams, or
 * concurrent graph branches) RACE -- serialise them on one stream.  Nothing in the array needs clearing between calls. */
size_t mrisr_ssim_static_scratch_bytes(void);
int mrisr_ssim_forward_static(const float* img1, const float* img2, float* per_sample, int N, int H, int W,
                              int window_size, float sigma, float val_range, void* stream);

/* ---- VGG perceptual term (replaces VGGFeatureExtractor / PerceptualLoss, utils/losses.py:83-151) ----------------
 * The feature network is a config list on the HOST: channels per 3x3 conv (+bias, +ReLU), -1 = MaxPool2d(2); the list must
 * end on a conv (vgg19.features[:36] = {64,64,-1,128,128,-1,256,256,256,256,-1,512,512,512,512,-1,512,512,512,512}).
 * `weights`: flat fp32 DEVICE arena (w0 [C0][3][3][3], b0 [C0], w1 [C1][C0][3][3], b1, ...) of mrisr_vgg_weight_elems()
 * elements, i.e. torchvision's conv weights in order; they are frozen (utils/losses.py:98-99).  `packed`: bf16 operand copies
 * made once by mrisr_vgg_pack_weights.  Images are N x 1 x H x W fp32 in [0,1]; the gray -> 3 channel repeat and the ImageNet
 * normalisation (utils/losses.py:107-114) happen inside.  bf16 operands, fp32 accumulation (tcgen05); H, W multiples of
 * 2^pools.  The workspace must be zero-filled once before its first use with a shape (zero-border invariant). */
int64_t mrisr_vgg_weight_elems(const int* host_cfg, int ncfg);
size_t  mrisr_vgg_packed_bytes(const int* host_cfg, int ncfg);
int     mrisr_vgg_pack_weights(const float* weights, const int* host_cfg, int ncfg, void* packed, size_t packed_bytes, void* stream);
size_t  mrisr_vgg_workspace_bytes(int N, int H, int W, const int* host_cfg, int ncfg, int grad);
/* feat: N x C_last x H/2^pools x W/2^pools fp32 (VGGFeatureExtractor.forward) */
int     mrisr_vgg_features(const void* packed, const float* weights, const float* img, float* feat, void* workspace,
                           size_t workspace_bytes, int N, int H, int W, const int* host_cfg, int ncfg, void* stream);
/* out1[0] = mean |f(pred) - f(target)| (loss_type 1) or mean (f(pred) - f(target))^2 (loss_type 2).
 * dpred_accum (nullable): dpred_accum += grad_scale * d(out1)/d(pred), N x 1 x H x W fp32; NULL skips the backward chain and
 * the workspace may then be sized with grad = 0. */
int     mrisr_vgg_perceptual_fwd_bwd(const void* packed, const float* weights, const float* pred, const float* target,
                                     float* dpred_accum, float* out1, void* workspace, size_t workspace_bytes, int N, int H,
                                     int W, const int* host_cfg, int ncfg, int loss_type, float grad_scale, void* stream);

/* fp32 parity mode of the same term: fp32 activations and weights on CUDA cores (no `packed` copies; the weights are re-laid
 * out inside the workspace on every call), for checking L6 / L7 against a float32 framework at 1e-4 the way mode MRISR_FP32
 * does for the U-Net.  Same arguments and results as the calls above; sized for test images (a 512^2 slice costs 0.2 TFLOP
 * per forward without tensor cores).  Workspace: zero-filled once per shape, from mrisr_vgg_workspace_bytes_fp32. */
size_t  mrisr_vgg_workspace_bytes_fp32(int N, int H, int W, const int* host_cfg, int ncfg, int grad);
int     mrisr_vgg_features_fp32(const float* weights, const float* img, float* feat, void* workspace, size_t workspace_bytes,
                                int N, int H, int W, const int* host_cfg, int ncfg, void* stream);
int     mrisr_vgg_perceptual_fwd_bwd_fp32(const float* weights, const float* pred, const float* target, float* dpred_accum,
                                          float* out1, void* workspace, size_t workspace_bytes, int N, int H, int W,
                                          const int* host_cfg, int ncfg, int loss_type, float grad_scale, void* stream);

/* ---- fused Adam (replaces torch.optim.Adam(lr, weight_decay) of scripts/train.py:186,318) -- */
/* L2 weight decay folded into the gradient (torch.optim.Adam semantics, not AdamW);
 * grad_scale multiplies g first (1/world_size for data parallel, 1/loss_scale for AMP). */
int mrisr_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                    float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                    float grad_scale, void* stream);
/* Same update with the step number AND the learning rate kept on the device, for CUDA-graph capture of the whole training
 * step (scripts/train.py:296-323 as one graph launch, with ReduceLROnPlateau of :189,397 still taking effect):
 * `state32` = 32 device bytes: int64 step count (advanced by this call), two floats of scratch, float lr (written by the
 * host whenever the scheduler changes it), 12 bytes reserved.  Two launches (advance + update); nothing step- or
 * lr-dependent is baked into the launch arguments. */
int mrisr_adam_step_dev(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                        float beta1, float beta2, float eps, float weight_decay, void* state32,
                        float grad_scale, void* stream);

/* ---- on-device input pipeline: dense uint8 slices [N][H][W] in, float32 N x 1 x H x W out ----------
 * Byte / index arithmetic reproduced bit for bit against the reference's PIL / numpy code (tests/test_gpu_preproc.py). */
/* transforms.ToTensor() of utils/dataset.py:124-125: dst = src / 255 (float32 division) */
int mrisr_u8_to_float(const uint8_t* src, float* dst, int64_t count, void* stream);
/* preprocess_image of scripts/infer.py:97-130: per slice min = np.percentile(x, lower), max = np.percentile(x, upper) (numpy
 * 'linear' method in float32, from a 256-bin histogram), clip, (x - min) / (max - min) when max > min (else the clipped raw
 * values, as the reference leaves them).  minmax: N x 2 floats (device), receives (min, max) per slice.  H * W < 2^24. */
int mrisr_percentile_normalize_u8(const uint8_t* src, float* dst, float* minmax, int N, int H, int W, float lower_percentile,
                                  float upper_percentile, void* stream);
/* augment_pair of utils/dataset.py:138-175 applied to one tensor of the pair (call once for the HR batch, once for the LR batch,
 * with the same per-sample decisions): hflip -> rotate (Pillow NEAREST, 16.16 fixed point; fill = int(mean(x/255) * 255)) ->
 * brightness -> contrast (Pillow ImageEnhance blends, truncating) -> additive noise (LR only) -> ToTensor.
 * params: N x MRISR_AUG_PARAM_INTS int32 (device): [0] flip, [1] rotate on, [2..7] Pillow's affine matrix in 16.16 fixed point
 *   (a0 a1 a2 / a3 a4 a5 with the half-pixel offset folded into a2, a5: utils/preprocess.py:pil_rotate_matrix), [8] brightness on,
 *   [9] factor (float bits), [10] contrast on, [11] factor (float bits), [12] noise on, [13] noise std in gray levels (float
 *   bits), [14] [15] Philox key.  The random DECISIONS are drawn on the host (a handful of scalars per sample, utils/dataset.py
 *   draws them with Python's `random`); the pixels never leave the device.
 * noise (nullable): N x H x W doubles added where noise is on (parity with np.random.normal arrays); NULL = Philox-4x32-10 +
 *   Box-Muller on the device.  tmp: N*H*W bytes, sums: 2*N uint64 (device scratch).  dst (float32) and / or dst_u8 receive the
 *   result.  H, W < 16384. */
#define MRISR_AUG_PARAM_INTS 16
int mrisr_augment_u8(const uint8_t* src, float* dst, uint8_t* dst_u8, uint8_t* tmp, uint64_t* sums, const int32_t* params,
                     const double* noise, int N, int H, int W, void* stream);

/* ---- batched evaluation ------------------------------------------------------------------------------
 * calculate_metrics of scripts/test_comparison.py:164-202 without its SSIM (mrisr_ssim_forward): out4[n] = (MSE, RMSE, MAE,
 * PSNR) of image n against reference n, both N x H x W float32; differences in float32 as the reference takes them, sums in
 * float64; PSNR = 100 when MSE < 1e-10, else 10 log10(1 / MSE) (data_range 1). */
size_t mrisr_image_metrics_scratch_bytes(int N, int H, int W);
int mrisr_image_metrics(const float* reference, const float* image, float* out4, void* scratch, int N, int H, int W, void* stream);
/* upscale_with_interpolation of scripts/test_comparison.py:92-134: cv2.resize(img, (2W, 2H)) on uint8 with method 0 = INTER_LINEAR,
 * 1 = INTER_LINEAR followed by filter2D([[-1,-1,-1],[-1,9,-1],[-1,-1,-1]]) ("sharp_bilinear"), 2 = INTER_CUBIC; dst = result / 255
 * (float32, N x 2H x 2W) and / or dst_u8.  Bit-exact against cv2 4.13's x86 build (fixed-point coefficients of 1/2048; the
 * cubic vertical pass in float32 for groups of four columns, as its SIMD body does).  tmp: N*2H*2W bytes. */
int mrisr_upscale2x_u8(const uint8_t* src, float* dst, uint8_t* dst_u8, uint8_t* tmp, int N, int H, int W, int method, void* stream);
/* skimage.exposure.match_histograms(source[n], reference[n]) of scripts/infer.py:303-307 for N image pairs (float32; the
 * reference image may have another size): out = interp(cdf_source(value), cdf_reference, reference values) in float64,
 * stored as float32 (the caller clips to [0, 1], infer.py:309).  Sorting is cub::DeviceRadixSort inside caller scratch. */
size_t mrisr_match_histograms_scratch_bytes(int N, int H, int W, int Href, int Wref);
int mrisr_match_histograms(const float* source, const float* reference, float* out, void* scratch, size_t scratch_bytes, int N, int H,
                           int W, int Href, int Wref, void* stream);

/* ---- k-space low-field simulation ---------------------------------------------------------------------
 * simulate_low_field_mri of utils/preprocessing.py:225-293 for N slices (float32 in [0,1], N x H x W): fft2 (single precision,
 * as numpy 2.x transforms a float32 array) -> keep the centre int(H*f) x int(W*f) block of the shifted spectrum -> add complex
 * Gaussian noise of standard deviation (noise_std / 255) * sqrt(H*W) / 10 (:271) -> ifft2 (complex128) -> magnitude ->
 * min-max rescaled to the input slice's own range.  out: N x H x W float64.  noise_real / noise_imag (both or neither):
 * N x H x W float64 arrays on the SHIFTED grid, exactly what np.random.normal would have produced (parity); NULL = Philox
 * + Box-Muller on the device keyed by `seed`.  cuFFT plans are cached per (device, N, H, W) and work inside `scratch`. */
size_t mrisr_kspace_scratch_bytes(int N, int H, int W);
int mrisr_simulate_low_field(const float* data, double* out, const double* noise_real, const double* noise_imag, void* scratch,
                             size_t scratch_bytes, int N, int H, int W, double kspace_crop_factor, double noise_std,
                             uint64_t seed, void* stream);
/* utils/extraction_utils.py:143-165: [clip(x, 0, 1) ->] cv2.resize(INTER_AREA) to exactly half size on float64 (H, W even)
 * [-> clip(x * 255, 0, 255) truncated to uint8, the PNG the reference writes].  dst (float64) and / or dst_u8: N x H/2 x W/2. */
int mrisr_area_half_f64(const double* src, double* dst, uint8_t* dst_u8, int N, int H, int W, int clip01, void* stream);
/* robust_normalize of utils/preprocessing.py:126-163 on N float64 slices: np.percentile(lower / upper) (float64, order
 * statistics from a segmented sort in scratch), clip, min-max to [target_lo, target_hi], float32 result; constant slices give
 * zeros.  window (nullable): N x 2 float64 (lower, upper). */
size_t mrisr_robust_normalize_scratch_bytes(int N, int H, int W);
int mrisr_robust_normalize_f64(const double* src, float* dst, double* window, void* scratch, size_t scratch_bytes, int N, int H, int W,
                               double lower_percentile, double upper_percentile, double target_lo, double target_hi, void* stream);

/* ==================================================================================================================
 * MRISR_INTERNAL -- test, diagnosis and measurement aids.  Exported so that tests/ and bench.py can reach single kernels and
 * internal tensors through the same C ABI; NOT part of the product surface INTEGRATION.md describes, and free to change.
 * Define MRISR_NO_INTERNAL before including this header to hide the declarations.
 * ================================================================================================================== */
#ifndef MRISR_NO_INTERNAL

/* Diagnostic (host-only arithmetic, no device needed): which kernel family runs convolution `conv_index` (0..19, the order of
 * models/unet_model.py:137-173 with the three up-path 1x1 convs at 8, 11, 14) in bf16 mode at input size H x W:
 * 0 = CUDA cores, 1 = NHWC TMA / tcgen05 (conv_tc, wgrad_tc), 2 = planar-8 sliding-window tcgen05 (conv_ring, wgrad_rows).
 * Levels <= 0 use 2 while three padded pixel rows of every 3x3 layer of the level fit in shared memory, else the whole
 * level falls back to 1 (512 x 512 input, BASELINE.json config 4).  -1 = bad argument. */
int mrisr_unet_layer_engine(int base_filters, int conv_index, int H, int W);

/* ---- convolution building block over padded NHWC bf16 tensors ---------------------------------
 * The implicit-GEMM kernels behind every nn.Conv2d of models/unet_model.py:29,34,72,101,152,168 (and of the VGG19
 * feature stack, utils/losses.py:90-95), exposed one layer at a time: the parity tests call them with both engines.
 * Tensors are [N][H+2][W+2][C] bf16 with a zero one-pixel border (pointer = padded origin); `w` is the reference's fp32
 * [Cout][Cin][ks][ks] weight.  engine: 0 = CUDA cores, 1 = tcgen05/TMEM/TMA (NHWC), 2 = tcgen05 planar-8 (below).  transposed != 0 computes the data
 * gradient instead: src1 = dL/d(out) with C1 = forward Cout channels, dst = dL/d(in) with `Cout` = forward Cin channels,
 * `w` still the forward weight [C1][Cout][ks][ks].  stats (nullable): [N][8][2] GroupNorm(8) (mean, rstd) of dst.
 * Border pixels of dst are never written (they must already be zero). */
/* engine 2 = tcgen05 sliding-window kernels over the planar-8 layout (narrow high-resolution layers): every tensor pointer
 * is then the START of a planar allocation of C/8 planes x plane_pixels x 8 channels, in which padded-linear pixel 0 sits
 * guard_pixels into each plane (mrisr_planar_geometry; guard contents are read but never influence stored outputs). */
int mrisr_planar_geometry(int N, int H, int W, long long* host_guard_pixels, long long* host_plane_pixels);
size_t mrisr_conv2d_scratch_bytes(int N, int H, int W, int C1, int C2, int Cout, int ks);
int mrisr_conv2d_bf16(const void* src1, int C1, const void* src2, int C2, const float* w, const float* bias, void* dst,
                      float* stats, void* scratch, size_t scratch_bytes, int N, int H, int W, int Cout, int ks,
                      int transposed, int engine, void* stream);
/* engine 3 (weight gradient only): row-structured tcgen05 kernel over the same planar-8 allocations as engine 2.
 * dw: fp32 [Cout][C1+C2][ks][ks], overwritten with sum over pixels of in (x) dy; dy must have a zero border (and, for
 * engine 2, zero guard pixels: the reduction runs over whole 128-pixel blocks; the guards of src1/src2 must be finite). */
int mrisr_conv2d_wgrad_bf16(const void* src1, int C1, const void* src2, int C2, const void* dy, float* dw, void* scratch,
                            size_t scratch_bytes, int N, int H, int W, int Cout, int ks, int engine, void* stream);

/* ---- debugging aid: copy one internal workspace tensor out as dense NCHW fp32 --------------- */
/* kind: 0 raw conv output R[l], 1 activation A[l], 2 pooled P[l], 3 low-res 1x1 output Q[l], 4 dR[l], 5 G[l]
 * (gradient wrt the input of conv l), 6 dQ[l], 7 UB, 8 XM (blend), 9 DY3.  dims_host[4] (HOST) receives N,C,H,W;
 * `out` may be NULL to query dims only.  Used by tests to localise a parity failure to one layer. */
int mrisr_unet_debug_tensor(void* workspace, int N, int H, int W, int base_filters, int precision, int kind,
                            int layer, float* out, int* dims_host, void* stream);

/* ---- measurement aids (used by bench.py; no reference counterpart) --------------------------- */
/* number of kernel launches issued by this library in this process so far */
long long mrisr_launch_count(void);
/* per-kernel-family CUDA-event timing on the launching stream: enable, run steps, then report.
 * mrisr_prof_report synchronises the device and writes "<tag> <calls> <total_ms>\n" lines (sorted by time)
 * into the HOST buffer buf (NUL-terminated, at most cap bytes); returns the number of tags. */
int mrisr_prof_enable(int on);
int mrisr_prof_report(char* host_buf, size_t cap);

#endif /* MRISR_NO_INTERNAL */

#ifdef __cplusplus
}
#endif
#endif /* MRISR_H */
