// Host-side launchers of the mrisr kernels (internal; the public boundary is include/mrisr.h).
#pragma once
#include "common.cuh"

namespace mrisr {

// ---------------------------------------------------------------- convolution (conv_generic.cu)
template <typename T>
struct ConvArgs {
    Act<T> src1;          // first C1 input channels
    Act<T> src2;          // next C2 input channels (skip-concat fused into the load), p == nullptr if unused
    const T* w;           // packed [ks*ks][Cin][Cout]
    const float* bias;    // [Cout] or nullptr
    Act<T> dst;           // C == Cout
    float* stat_part;     // [N][slots][8][2] partial (sum, sumsq) per GroupNorm group, or nullptr
    int ks;               // 1 or 3
    bool relu = false;    // conv_tc only: ReLU in the epilogue (VGG conv + ReLU)
    const T* mask = nullptr;   // conv_tc only: zero the output where mask <= 0 (same layout as dst; ReLU backward in dgrad)
};
// slots written per sample by launch_conv_generic for this shape
int conv_generic_stat_slots(int H, int W, int Cout);
template <typename T> int launch_conv_generic(const ConvArgs<T>& a, cudaStream_t s);

template <typename T>
struct WgradArgs {
    Act<T> src1, src2;    // forward inputs (concat along channels)
    Act<T> dy;            // gradient wrt the raw conv output (zero borders)
    float* dw;            // [Cout][Cin][ks][ks] fp32 (reference NCHW weight layout), overwritten
    float* partial;       // scratch, >= wgrad_generic_scratch_elems() floats
    int ks;
};
size_t wgrad_generic_scratch_elems(int N, int H, int W, int Cin, int Cout, int ks);
template <typename T> int launch_wgrad_generic(const WgradArgs<T>& a, cudaStream_t s);

// fp32 NCHW [Cout][Cin][ks][ks] -> forward pack [tap][Cin][Cout] and dgrad pack [tap'][Cout][Cin] (tap' = flipped)
template <typename T>
int launch_pack_weights(const float* w, T* w_fwd, T* w_dgrad, int Cout, int Cin, int ks, cudaStream_t s);

// ---------------------------------------------------------------- first layer, 1 -> f channels (conv_first.cu), bf16 + planar output
// ConvArgs::w is the generic pack [tap][1][Cout]; stat_part receives [N][conv_first_stat_slots][8][2] for launch_gn_finalize
bool conv_first_supported(const Act<__nv_bfloat16>& src, const Act<__nv_bfloat16>& dst, int ks);
int conv_first_stat_slots(int H, int W, int Cout);
size_t wgrad_first_scratch_elems(int N, int H, int W, int Cout);
int launch_conv_first(const ConvArgs<__nv_bfloat16>& a, cudaStream_t s);
int launch_wgrad_first(const WgradArgs<__nv_bfloat16>& a, cudaStream_t s);

// ---------------------------------------------------------------- tcgen05 convolution (conv_tc.cu, wgrad_tc.cu), bf16 only
// ConvArgs::w is the K-major pack of launch_pack_weights_tc ([tap][Cout][Cin] forward, [tap'][Cin][Cout] dgrad);
// ConvArgs::stat_part receives [tiles][2][8][2] partials (conv_tc_stat_elems floats) consumed by launch_gn_finalize_tc.
bool conv_tc_supported(int H, int W, int C1, int C2, int Cout, int ks);
int conv_tc_stat_elems(int N, int H, int W);
int launch_conv_tc(const ConvArgs<__nv_bfloat16>& a, cudaStream_t s);
int launch_gn_finalize_tc(const float* part, float* stats, int N, int H, int W, double count, cudaStream_t s);
int launch_pack_weights_tc(const float* w, __nv_bfloat16* wf, __nv_bfloat16* wd, int Cout, int Cin, int ks, cudaStream_t s);
// planar-8 sliding-window kernels for narrow high-resolution layers (conv_ring.cu); weights packed by launch_pack_weights_ring
bool conv_ring_supported(int H, int W, int C1, int C2, int Cout, int ks);
int launch_conv_ring(const ConvArgs<__nv_bfloat16>& a, cudaStream_t s);
// launch_conv_ring writes per-warp statistics records (conv_ring_stat_elems floats) into ConvArgs::stat_part;
// launch_gn_finalize_ring turns them into [N][8][2] (mean, rstd)
int conv_ring_stat_elems(int N, int H, int W, int Cin, int Cout);
int launch_gn_finalize_ring(const float* rec, float* stats, int N, int H, int W, int Cin, int Cout, double count, cudaStream_t s);
int launch_pack_weights_ring(const float* w, __nv_bfloat16* wf, __nv_bfloat16* wd, int Cout, int Cin, int ks, cudaStream_t s);
bool wgrad_ring_supported(int H, int W, int C1, int C2, int Cout, int ks);
size_t wgrad_ring_scratch_elems(int N, int H, int W, int Cin, int Cout, int ks);
int launch_wgrad_ring(const WgradArgs<__nv_bfloat16>& a, cudaStream_t s);
// row-structured weight gradient (wgrad_rows.cu): vertical taps folded into M; preferred over wgrad_ring when supported
bool wgrad_rows_supported(int N, int H, int W, int C1, int C2, int Cout, int ks);
size_t wgrad_rows_scratch_elems(int N, int H, int W, int Cin, int Cout);
int launch_wgrad_rows(const WgradArgs<__nv_bfloat16>& a, cudaStream_t s);
bool wgrad_tc_supported(int H, int W, int C1, int C2, int Cout, int ks);
size_t wgrad_tc_scratch_elems(int N, int H, int W, int C1, int C2, int Cout, int ks);
int launch_wgrad_tc(const WgradArgs<__nv_bfloat16>& a, cudaStream_t s);
// wgrad_tc3.cu: horizontal taps merged into one fetched tile (64-channel chunks, Cout 64 / 128); launch_wgrad_tc dispatches
bool wgrad_tc3_supported(int N, int H, int W, int C1, int C2, int Cout, int ks);
size_t wgrad_tc3_scratch_elems(int N, int H, int W, int C1, int C2, int Cout);
int launch_wgrad_tc3(const WgradArgs<__nv_bfloat16>& a, cudaStream_t s);

// ---------------------------------------------------------------- normalisation / activation (norm_act.cu)
template <typename T> int launch_pack_input(const float* x, Act<T> dst, cudaStream_t s);

int stat_slots_generic(int H, int W, int C);
// partial (sum, sumsq) per (n, slot, group) of an activation tensor
template <typename T> int launch_gn_stats(Act<T> r, float* part, cudaStream_t s);
// part [N][slots][8][2] -> stats [N][8][2] = (mean, rstd); count = elements per group
int launch_gn_finalize(const float* part, float* stats, int N, int slots, double count, cudaStream_t s);

// a = lrelu(gn(r));  pooled (optional, p != nullptr) = maxpool2x2(a)
template <typename T>
int launch_gn_apply(Act<T> r, const float* stats, const float* gamma, const float* beta, Act<T> a, Act<T> pooled,
                    cudaStream_t s);

// dst[2h x 2w] = bilinear x2 (align_corners=True) of src[h x w]
template <typename T> int launch_upsample2x(Act<T> src, Act<T> dst, cudaStream_t s);
constexpr int MRISR_NOT_APPLICABLE = -12345;      // internal: "this fused form does not cover the shape, use the plain path"
// same, also leaving dst's GroupNorm partials (gn_stats layout, *slots per sample) in `part`; MRISR_NOT_APPLICABLE if the shape needs the plain path
template <typename T> int launch_upsample2x_stats(Act<T> src, Act<T> dst, float* part, size_t part_cap_floats, int* slots, cudaStream_t s);
// dsrc = adjoint(upsample)(ddst) (+ addend if addend.p != nullptr)
template <typename T> int launch_upsample2x_bwd(Act<T> ddst, Act<T> addend, Act<T> dsrc, cudaStream_t s);

// xm = sig(alpha) * lrelu(gn(r_bil)) + (1 - sig(alpha)) * lrelu(gn(pixel_shuffle(r_ps)))
template <typename T>
struct BlendArgs {
    Act<T> r_bil;  const float* st_bil; const float* g_bil; const float* b_bil;     // [N,2H,2W,Ch]
    Act<T> r_ps;   const float* st_ps;  const float* g_ps;  const float* b_ps;      // [N,H,W,4Ch] unshuffled
    const float* alpha;                                                             // scalar parameter
    Act<T> xm;                                                                      // [N,2H,2W,Ch]
};
template <typename T> int launch_head_blend(const BlendArgs<T>& a, cudaStream_t s);
// d(alpha) = sigmoid'(alpha) * sum( dxm * (xb - xp) ) from the per-channel sums  sum g * xb,  sum g * xp  that the two
// blend-kind gn_bwd launches write ([2][nch])
int launch_dalpha_from_channels(const float* sums, int nch, const float* alpha, float* dalpha, cudaStream_t s);

// y = sigmoid( w1x1 . lrelu(gn(r)) + b )      (final_conv.1..3 + sigmoid, models/unet_model.py:169-172,211)
template <typename T>
int launch_final_head(Act<T> r, const float* stats, const float* gamma, const float* beta, const float* w1x1,
                      const float* b1x1, float* y, cudaStream_t s);

// ---- GroupNorm + LeakyReLU backward -------------------------------------------------------
// Gradient wrt the post-activation tensor is described by a "source":
enum GradSrcKind : int {
    SRC_DIRECT = 0,       // g[n,y,x,coff+c]
    SRC_DIRECT_POOL = 1,  // g[n,y,x,coff+c] + (argmax of the 2x2 window ? gpool[n,y/2,x/2,c] : 0)
    SRC_POOL_ONLY = 2,    // only the pool term (unused by the U-Net, kept for VGG)
    SRC_OUTER = 3,        // dl[n,y,x] * wvec[c],  dl = dy*y*(1-y)   (final 1x1 conv + sigmoid)
    SRC_BLEND_BIL = 4,    // sig(alpha) * g[n,y,x,c]
    SRC_BLEND_PS = 5,     // (1-sig(alpha)) * g[n,2y+i,2x+j,c/4], c = 4c'+2i+j   (pixel-shuffle adjoint)
};
template <typename T>
struct GnBwdArgs {
    int kind;
    Act<T> r;  const float* stats; const float* gamma; const float* beta;   // gamma/beta indexed c (or c/4 for SRC_BLEND_PS)
    Act<T> g;  int coff;          // main gradient tensor and channel offset into it
    Act<T> gpool;                 // gradient wrt the pooled tensor (SRC_DIRECT_POOL)
    const float* dy; const float* y; const float* wvec;   // SRC_OUTER (dy, y: [N,H,W] fp32)
    const float* alpha;           // SRC_BLEND_*
    float* part;                  // scratch [N][slots][C][2]
    float* mstat;                 // scratch [N][8][2]  (m1, m2)
    float* dgamma; float* dbeta;  // outputs [C] (or [C/4] for SRC_BLEND_PS), overwritten
    float* dwvec;  float* dbias;  // SRC_OUTER only: d(w1x1) [C], d(b1x1) [1]
    Act<T> dr;                    // output: gradient wrt the raw conv output
};
int gn_bwd_slots(int H, int W, int C);
template <typename T> int launch_gn_bwd(const GnBwdArgs<T>& a, cudaStream_t s);

// per-channel sum over all pixels (conv bias gradient): out[c] = sum_{n,y,x} t[n,y,x,c]
size_t colsum_scratch_elems(int N, int H, int W, int C);
template <typename T> int launch_colsum(Act<T> t, float* out, float* scratch, cudaStream_t s);

// ---------------------------------------------------------------- loss / optimiser (loss.cu, adam.cu)
size_t l1_ssim_scratch_bytes(int N, int H, int W);
int launch_l1_ssim(const float* pred, const float* target, float* dpred, float* out5, void* scratch, int N, int H,
                   int W, float l1_w, float ssim_w, int win, float sigma, float val_range, cudaStream_t s);
int launch_ssim_forward(const float* a, const float* b, float* per_sample, void* scratch, int N, int H, int W,
                        int win, float sigma, float val_range, cudaStream_t s);
int launch_adam(float* p, const float* g, float* m, float* v, int64_t n, float lr, float b1, float b2, float eps,
                float wd, int step, float grad_scale, cudaStream_t s);
int launch_adam_dev(float* p, const float* g, float* m, float* v, int64_t n, float b1, float b2, float eps, float wd,
                    void* state32, float grad_scale, cudaStream_t s);

}  // namespace mrisr
