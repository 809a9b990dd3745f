// VGG perceptual term of CombinedLoss on the tcgen05 convolution kernels (conv_tc.cu):
//   VGGFeatureExtractor.forward  utils/losses.py:105-118  (gray -> 3 channels, ImageNet normalisation, vgg19.features[:36])
//   PerceptualLoss.forward       utils/losses.py:138-151  (L1 / MSE between the features of generated and target,
//                                                          target branch without gradient)
// and the data gradient of the generated branch back to the image (autograd in the reference; the VGG weights are frozen,
// utils/losses.py:98-99, so there is no weight gradient).  The network is described by a config list (channels per conv,
// -1 = MaxPool2d(2)), so VGG16 or a shorter prefix is a config; VGG19[:36] is the parity target.
//
// Layout: padded NHWC bf16 with a zero border (zero padding of the NORMALISED image, as the reference pads).
// First conv (3 -> 64, K = 27): the image is packed as its own 3x3 patches, [pixel][tap * 3 + channel] padded to 32 columns
// (out-of-image taps are the zero padding of the normalised image), so the layer is ONE K = 32 GEMM step per 128-pixel
// tile on the 1x1 path of the implicit-GEMM kernel instead of nine K = 16 steps with 13 zero channels each
// (3.28 -> see profiles/ ms per call at 64 x 512^2).  Its data gradient keeps the 16-channel form (3 real channels).
// Fusions: bias + ReLU in the conv epilogue; ReLU backward (mask by the saved activation) in the dgrad epilogue;
// max-pool backward routes to the first maximum in scan order and applies the ReLU mask in one pass.
#include "../../include/mrisr.h"
#include "kernels.cuh"

namespace mrisr {
namespace {
using bf16 = __nv_bfloat16;
constexpr int NTHR = 256;
constexpr int kMaxCfg = 64;
constexpr int kCin0 = 16;                 // padded input channels of the first conv's data gradient
constexpr int kK0 = 32;                   // patch columns of the first conv's forward operand: 9 taps x 3 channels, padded

__constant__ float c_vgg_mean[3] = {0.485f, 0.456f, 0.406f};     // utils/losses.py:7
__constant__ float c_vgg_std[3] = {0.229f, 0.224f, 0.225f};      // utils/losses.py:8

inline int grid_for(int64_t total) {
    int64_t b = ceil_div64(total, NTHR);
    if (b > 148 * 16) b = 148 * 16;
    if (b < 1) b = 1;
    return (int)b;
}

// x: N x 1 x H x W fp32 -> [N][H+2][W+2][32] bf16 patches: column tap * 3 + c = (x[y + ty - 1, x + tx - 1] - mean_c) / std_c
// inside the image, 0 outside (repeat + normalise + the conv's zero padding, utils/losses.py:107-114); columns 27..31 zero
__global__ void __launch_bounds__(NTHR) vgg_pack_input_kernel(const float* __restrict__ x, Act<bf16> dst) {
    const int64_t total = (int64_t)dst.N * dst.H * dst.W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int xx = (int)(i % dst.W), yy = (int)((i / dst.W) % dst.H), n = (int)(i / ((int64_t)dst.W * dst.H));
        const float* img = x + (int64_t)n * dst.H * dst.W;
        float o[kK0];
#pragma unroll
        for (int k = 27; k < kK0; ++k) o[k] = 0.f;
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
            const int sy = yy + tap / 3 - 1, sx = xx + tap % 3 - 1;
            const bool in = sy >= 0 && sy < dst.H && sx >= 0 && sx < dst.W;
            const float v = in ? __ldg(img + (int64_t)sy * dst.W + sx) : 0.f;
#pragma unroll
            for (int c = 0; c < 3; ++c) o[tap * 3 + c] = in ? (v - c_vgg_mean[c]) / c_vgg_std[c] : 0.f;
        }
        bf16* p = dst.at(n, yy, xx, 0);           // 64 B per pixel as two full-sector stores
#pragma unroll
        for (int k = 0; k < kK0; k += 16)
            st_global_256(p + k, pack_bf16x2(o[k], o[k + 1]), pack_bf16x2(o[k + 2], o[k + 3]), pack_bf16x2(o[k + 4], o[k + 5]),
                          pack_bf16x2(o[k + 6], o[k + 7]), pack_bf16x2(o[k + 8], o[k + 9]), pack_bf16x2(o[k + 10], o[k + 11]),
                          pack_bf16x2(o[k + 12], o[k + 13]), pack_bf16x2(o[k + 14], o[k + 15]));
    }
}

// dimg[n,y,x] += scale * sum_c g[n,y,x,c] / std_c      (adjoint of repeat + normalise)
template <typename T>
__global__ void __launch_bounds__(NTHR) vgg_dimg_kernel(Act<T> g, float* __restrict__ dimg, float scale) {
    const int64_t total = (int64_t)g.N * g.H * g.W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int xx = (int)(i % g.W), yy = (int)((i / g.W) % g.H), n = (int)(i / ((int64_t)g.W * g.H));
        float v[8];
        load8(g.at(n, yy, xx, 0), v);
        dimg[i] += scale * (v[0] / c_vgg_std[0] + v[1] / c_vgg_std[1] + v[2] / c_vgg_std[2]);
    }
}

template <typename T>
__global__ void __launch_bounds__(NTHR) vgg_maxpool_kernel(Act<T> a, Act<T> p) {
    const int nv = a.C / 8;
    const int64_t total = (int64_t)p.N * p.H * p.W * nv;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int cv = (int)(i % nv);
        const int x = (int)((i / nv) % p.W), y = (int)((i / ((int64_t)nv * p.W)) % p.H);
        const int n = (int)(i / ((int64_t)nv * p.W * p.H));
        float m[8];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float v[8];
            load8(a.at(n, 2 * y + (q >> 1), 2 * x + (q & 1), cv * 8), v);
#pragma unroll
            for (int k = 0; k < 8; ++k) m[k] = q == 0 ? v[k] : fmaxf(m[k], v[k]);
        }
        store8(p.at(n, y, x, cv * 8), m);
    }
}

// da[pos] = (pos is the first maximum of its 2x2 window in scan order, and a[pos] > 0) ? dp : 0
// (max_pool2d backward followed by the ReLU backward of the conv that produced a)
template <typename T>
__global__ void __launch_bounds__(NTHR) vgg_pool_bwd_kernel(Act<T> a, Act<T> dp, Act<T> da) {
    const int nv = a.C / 8;
    const int64_t total = (int64_t)dp.N * dp.H * dp.W * nv;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int cv = (int)(i % nv);
        const int x = (int)((i / nv) % dp.W), y = (int)((i / ((int64_t)nv * dp.W)) % dp.H);
        const int n = (int)(i / ((int64_t)nv * dp.W * dp.H));
        float v[4][8], g[8];
#pragma unroll
        for (int q = 0; q < 4; ++q) load8(a.at(n, 2 * y + (q >> 1), 2 * x + (q & 1), cv * 8), v[q]);
        load8(dp.at(n, y, x, cv * 8), g);
        int best[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            best[k] = 0;
            float bv = v[0][k];
#pragma unroll
            for (int q = 1; q < 4; ++q)
                if (v[q][k] > bv) { bv = v[q][k]; best[k] = q; }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float o[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = (best[k] == q && v[q][k] > 0.f) ? g[k] : 0.f;
            store8(da.at(n, 2 * y + (q >> 1), 2 * x + (q & 1), cv * 8), o);
        }
    }
}

// Feature loss: partial[b] = sum |fg - ft| (L1) or sum (fg - ft)^2 (MSE); dfg = d(mean loss)/d(fg) masked by fg > 0
// (ReLU backward of the last conv).  loss_type: 1 = L1, 2 = MSE   (utils/losses.py:128-131)
template <typename T>
__global__ void __launch_bounds__(NTHR) vgg_feat_loss_kernel(Act<T> fg, Act<T> ft, Act<T> dfg, float* __restrict__ partial,
                                                             int loss_type, float inv_numel) {
    __shared__ float s_w[NTHR / 32];
    const int nv = fg.C / 8;
    const int64_t total = (int64_t)fg.N * fg.H * fg.W * nv;
    float acc = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int cv = (int)(i % nv);
        const int x = (int)((i / nv) % fg.W), y = (int)((i / ((int64_t)nv * fg.W)) % fg.H);
        const int n = (int)(i / ((int64_t)nv * fg.W * fg.H));
        float a[8], b[8], o[8];
        load8(fg.at(n, y, x, cv * 8), a);
        load8(ft.at(n, y, x, cv * 8), b);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float d = a[k] - b[k];
            float g;
            if (loss_type == 1) { acc += fabsf(d); g = d > 0.f ? inv_numel : (d < 0.f ? -inv_numel : 0.f); }
            else { acc = fmaf(d, d, acc); g = 2.f * d * inv_numel; }
            o[k] = a[k] > 0.f ? g : 0.f;
        }
        if (dfg.p) store8(dfg.at(n, y, x, cv * 8), o);
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int w = 0; w < NTHR / 32; ++w) t += s_w[w];
        partial[blockIdx.x] = t;
    }
}
__global__ void vgg_loss_finalize_kernel(const float* __restrict__ partial, int nblocks, double inv_numel, float* __restrict__ out1) {
    double t = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 32) t += (double)partial[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (threadIdx.x == 0) out1[0] = (float)(t * inv_numel);
}

// copy a padded NHWC bf16 tensor out as dense NCHW fp32 (test aid: VGGFeatureExtractor.forward parity)
template <typename T>
__global__ void vgg_unpack_kernel(Act<T> a, float* __restrict__ out) {
    const int64_t total = (int64_t)a.N * a.C * a.H * a.W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int x = (int)(i % a.W), y = (int)((i / a.W) % a.H);
        const int c = (int)((i / ((int64_t)a.W * a.H)) % a.C), n = (int)(i / ((int64_t)a.W * a.H * a.C));
        out[i] = to_f(*a.at(n, y, x, c));
    }
}

// first conv: fp32 [Cout][3][3][3] -> wf[Cout][32] (column tap * 3 + ci, 27..31 zero: the patch operand of
// vgg_pack_input_kernel) and the data-gradient form wd[8 - tap][16][Cout] (input channels 3..15 zero)
__global__ void vgg_pack_first_kernel(const float* __restrict__ w, bf16* __restrict__ wf, bf16* __restrict__ wd, int Cout) {
    const int total = 9 * Cout * kCin0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int ci = i % kCin0, co = (i / kCin0) % Cout, tap = i / (kCin0 * Cout);
        const bf16 v = __float2bfloat16_rn(ci < 3 ? w[((int64_t)co * 3 + ci) * 9 + tap] : 0.f);
        wd[((int64_t)(8 - tap) * kCin0 + ci) * Cout + co] = v;
        if (ci < 3) wf[(int64_t)co * kK0 + tap * 3 + ci] = v;
        if (tap == 0 && ci < kK0 - 27) wf[(int64_t)co * kK0 + 27 + ci] = __float2bfloat16_rn(0.f);
    }
}

struct Carve {
    char* base; size_t used = 0;
    explicit Carve(void* b) : base((char*)b) {}
    void* take(size_t bytes) { void* p = base ? base + used : nullptr; used += (bytes + 255) & ~(size_t)255; return p; }
};

struct VggNet {
    int n_items = 0, n_conv = 0, n_pool = 0;
    int item[kMaxCfg];             // channels, or -1 for a pool
    int cin[kMaxCfg], cout[kMaxCfg], lvl[kMaxCfg];     // per conv: channels and resolution level (H >> lvl)
    int pool_after[kMaxCfg];       // per conv: 1 if a pool follows it
    int64_t w_off[kMaxCfg], b_off[kMaxCfg];            // element offsets into the flat fp32 (w0, b0, w1, b1, ...) arena
    size_t wf_off[kMaxCfg], wd_off[kMaxCfg], packed_bytes = 0;
    bool ok = false;
    VggNet(const int* cfg, int ncfg) {
        if (!cfg || ncfg <= 0 || ncfg > kMaxCfg || cfg[0] <= 0) return;
        int c = 3, l = 0;
        int64_t off = 0;
        Carve pk(nullptr);
        for (int i = 0; i < ncfg; ++i) {
            item[i] = cfg[i];
            if (cfg[i] < 0) {
                if (n_conv == 0 || pool_after[n_conv - 1]) return;      // pool must follow a conv
                pool_after[n_conv - 1] = 1;
                ++n_pool; ++l;
            } else {
                if (cfg[i] % 16) return;
                cin[n_conv] = c; cout[n_conv] = cfg[i]; lvl[n_conv] = l; pool_after[n_conv] = 0;
                w_off[n_conv] = off; off += (int64_t)cfg[i] * c * 9;
                b_off[n_conv] = off; off += cfg[i];
                const int cpad = n_conv == 0 ? kCin0 : c;
                wf_off[n_conv] = pk.used; pk.take((size_t)9 * cfg[i] * cpad * sizeof(bf16));
                wd_off[n_conv] = pk.used; pk.take((size_t)9 * cfg[i] * cpad * sizeof(bf16));
                c = cfg[i];
                ++n_conv;
            }
        }
        n_items = ncfg;
        packed_bytes = pk.used;
        ok = n_conv > 0 && !pool_after[n_conv - 1];                     // features end on a conv + ReLU (idx 35 = relu5_4)
    }
    int cin_padded(int i) const { return i == 0 ? kCin0 : cin[i]; }
};

static Act<bf16> make_nhwc(Carve& ar, int N, int H, int W, int C) {
    Act<bf16> a;
    a.N = N; a.H = H; a.W = W; a.C = C; a.ps = C; a.cs = 8;
    a.p = (bf16*)ar.take((size_t)a.rows() * C * sizeof(bf16));
    return a;
}

struct VggWs {
    Act<bf16> X0g, X0t, Ag[kMaxCfg], At[kMaxCfg], Pg[kMaxCfg], Pt[kMaxCfg], dA[kMaxCfg], dP[kMaxCfg], dX0;
    float* partial;
    size_t bytes;
    VggWs(void* base, const VggNet& net, int N, int H, int W, bool grad) {
        Carve ar(base);
        X0g = make_nhwc(ar, N, H, W, kK0);
        X0t = make_nhwc(ar, N, H, W, kK0);
        for (int i = 0; i < net.n_conv; ++i) {
            const int h = H >> net.lvl[i], w = W >> net.lvl[i];
            Ag[i] = make_nhwc(ar, N, h, w, net.cout[i]);
            At[i] = make_nhwc(ar, N, h, w, net.cout[i]);
            if (net.pool_after[i]) { Pg[i] = make_nhwc(ar, N, h / 2, w / 2, net.cout[i]); Pt[i] = make_nhwc(ar, N, h / 2, w / 2, net.cout[i]); }
            if (grad) {
                dA[i] = make_nhwc(ar, N, h, w, net.cout[i]);
                if (net.pool_after[i]) dP[i] = make_nhwc(ar, N, h / 2, w / 2, net.cout[i]);
            }
        }
        if (grad) dX0 = make_nhwc(ar, N, H, W, kCin0);
        partial = (float*)ar.take((size_t)148 * 16 * sizeof(float));
        bytes = ar.used;
    }
};

#define TRY(expr) do { int _rc = (expr); if (_rc) return _rc; } while (0)

static int check_args(const VggNet& net, int N, int H, int W) {
    if (!net.ok) { set_error("vgg: bad config (channels must be multiples of 16, a pool must follow a conv, the list must end on a conv)"); return MRISR_E_BADARG; }
    if (N <= 0 || H <= 0 || W <= 0) { set_error("vgg: N, H, W must be positive"); return MRISR_E_BADARG; }
    const int div = 1 << net.n_pool;
    if (H % div || W % div) { set_error("vgg: H and W must be multiples of %d (%d poolings), got %dx%d", div, net.n_pool, H, W); return MRISR_E_SHAPE; }
    for (int i = 0; i < net.n_conv; ++i)
        if (!conv_tc_supported(H >> net.lvl[i], W >> net.lvl[i], i == 0 ? kK0 : net.cin[i], 0, net.cout[i], i == 0 ? 1 : 3) ||
            !conv_tc_supported(H >> net.lvl[i], W >> net.lvl[i], net.cout[i], 0, net.cin_padded(i), 3)) {
            set_error("vgg: conv %d (%d -> %d) is not supported by the tcgen05 kernel", i, net.cin[i], net.cout[i]);
            return MRISR_E_SHAPE;
        }
    return 0;
}

// one branch of the feature extractor; A / P receive every activation (saved for the backward of the generated branch)
static int vgg_features(const VggNet& net, const void* packed, const float* weights, const float* img, Act<bf16> X0,
                        Act<bf16>* A, Act<bf16>* P, cudaStream_t s) {
    {
        ProfScope prof_("vgg_pack_input", s);
        vgg_pack_input_kernel<<<grid_for((int64_t)X0.N * X0.H * X0.W), NTHR, 0, s>>>(img, X0);
        MRISR_LAUNCH_CHECK();
    }
    const Act<bf16> none{nullptr, 0, 0, 0, 0};
    Act<bf16> cur = X0;
    for (int i = 0; i < net.n_conv; ++i) {
        ConvArgs<bf16> a;
        a.src1 = cur; a.src2 = none; a.w = (const bf16*)((const char*)packed + net.wf_off[i]);
        a.bias = weights + net.b_off[i]; a.dst = A[i]; a.stat_part = nullptr; a.relu = true;
        a.ks = i == 0 ? 1 : 3;                       // the first layer's 3x3 window is already unrolled into its operand
        TRY(launch_conv_tc(a, s));
        cur = A[i];
        if (net.pool_after[i]) {
            ProfScope prof_("vgg_maxpool", s);
            vgg_maxpool_kernel<bf16><<<grid_for((int64_t)P[i].N * P[i].H * P[i].W * (P[i].C / 8)), NTHR, 0, s>>>(A[i], P[i]);
            MRISR_LAUNCH_CHECK();
            cur = P[i];
        }
    }
    return 0;
}


// ---- fp32 parity mode (CUDA cores): the same term with fp32 activations and weights on conv_generic<float>, so that the
// north_star's 1e-4 fp32-mode bound can be tested for L6 / L7 as it is for the U-Net.  Test-sized images only: a 512^2 image
// is 204 GFLOP per forward on CUDA cores.  Layout: padded NHWC fp32; the first conv's 3 input channels are padded to 8.
__global__ void __launch_bounds__(NTHR) vgg_pack_input_f32_kernel(const float* __restrict__ x, Act<float> dst) {
    const int64_t total = (int64_t)dst.N * dst.H * dst.W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int xx = (int)(i % dst.W), yy = (int)((i / dst.W) % dst.H), n = (int)(i / ((int64_t)dst.W * dst.H));
        const float v = x[i];
        float o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < 3; ++c) o[c] = (v - c_vgg_mean[c]) / c_vgg_std[c];
        store8(dst.at(n, yy, xx, 0), o);
    }
}
// [Cout][3][3][3] -> [Cout][8][3][3] with zero input channels 3..7
__global__ void vgg_expand_first_kernel(const float* __restrict__ w, float* __restrict__ w8, int Cout) {
    const int total = Cout * 8 * 9;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int tap = i % 9, ci = (i / 9) % 8, co = i / 72;
        w8[i] = ci < 3 ? w[((int64_t)co * 3 + ci) * 9 + tap] : 0.f;
    }
}
// in place over the interior of a padded tensor: relu (mask == nullptr) or zero where mask <= 0
__global__ void __launch_bounds__(NTHR) vgg_relu_mask_f32_kernel(Act<float> a, Act<float> mask) {
    const int nv = a.C / 8;
    const int64_t total = (int64_t)a.N * a.H * a.W * nv;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int cv = (int)(i % nv);
        const int x = (int)((i / nv) % a.W), y = (int)((i / ((int64_t)nv * a.W)) % a.H);
        const int n = (int)(i / ((int64_t)nv * a.W * a.H));
        float v[8], m[8];
        load8(a.at(n, y, x, cv * 8), v);
        if (mask.p) load8(mask.at(n, y, x, cv * 8), m);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = mask.p ? (m[k] > 0.f ? v[k] : 0.f) : fmaxf(v[k], 0.f);
        store8(a.at(n, y, x, cv * 8), v);
    }
}

static Act<float> make_nhwc_f32(Carve& ar, int N, int H, int W, int C) {
    Act<float> a;
    a.N = N; a.H = H; a.W = W; a.C = C; a.ps = C; a.cs = 8;
    a.p = (float*)ar.take((size_t)a.rows() * C * sizeof(float));
    return a;
}
struct VggWsF {
    Act<float> X0g, X0t, Ag[kMaxCfg], At[kMaxCfg], Pg[kMaxCfg], Pt[kMaxCfg], dA[kMaxCfg], dP[kMaxCfg], dX0;
    float* wf[kMaxCfg]; float* wd[kMaxCfg]; float* w8;
    float* partial;
    size_t bytes, zero_bytes;
    VggWsF(void* base, const VggNet& net, int N, int H, int W, bool grad) {
        Carve ar(base);
        X0g = make_nhwc_f32(ar, N, H, W, 8);
        X0t = make_nhwc_f32(ar, N, H, W, 8);
        for (int i = 0; i < net.n_conv; ++i) {
            const int h = H >> net.lvl[i], w = W >> net.lvl[i];
            Ag[i] = make_nhwc_f32(ar, N, h, w, net.cout[i]);
            At[i] = make_nhwc_f32(ar, N, h, w, net.cout[i]);
            if (net.pool_after[i]) { Pg[i] = make_nhwc_f32(ar, N, h / 2, w / 2, net.cout[i]); Pt[i] = make_nhwc_f32(ar, N, h / 2, w / 2, net.cout[i]); }
            if (grad) {
                dA[i] = make_nhwc_f32(ar, N, h, w, net.cout[i]);
                if (net.pool_after[i]) dP[i] = make_nhwc_f32(ar, N, h / 2, w / 2, net.cout[i]);
            }
        }
        if (grad) dX0 = make_nhwc_f32(ar, N, H, W, 8);
        zero_bytes = ar.used;                                   // everything above relies on the zero border
        for (int i = 0; i < net.n_conv; ++i) {
            const size_t n = (size_t)9 * net.cout[i] * (i == 0 ? 8 : net.cin[i]) * sizeof(float);
            wf[i] = (float*)ar.take(n);
            wd[i] = (float*)ar.take(n);
        }
        w8 = (float*)ar.take((size_t)net.cout[0] * 72 * sizeof(float));
        partial = (float*)ar.take((size_t)148 * 16 * sizeof(float));
        bytes = ar.used;
    }
};
static int check_args_f32(const VggNet& net, int N, int H, int W) {
    if (!net.ok) { set_error("vgg: bad config"); return MRISR_E_BADARG; }
    if (N <= 0 || H <= 0 || W <= 0) { set_error("vgg: N, H, W must be positive"); return MRISR_E_BADARG; }
    const int div = 1 << net.n_pool;
    if (H % div || W % div) { set_error("vgg: H and W must be multiples of %d (%d poolings), got %dx%d", div, net.n_pool, H, W); return MRISR_E_SHAPE; }
    return 0;
}
static int vgg_pack_f32(const VggNet& net, const float* weights, VggWsF& ws, cudaStream_t s) {
    vgg_expand_first_kernel<<<ceil_div(net.cout[0] * 72, 256), 256, 0, s>>>(weights + net.w_off[0], ws.w8, net.cout[0]);
    MRISR_LAUNCH_CHECK();
    for (int i = 0; i < net.n_conv; ++i)
        TRY(launch_pack_weights<float>(i == 0 ? ws.w8 : weights + net.w_off[i], ws.wf[i], ws.wd[i], net.cout[i], i == 0 ? 8 : net.cin[i], 3, s));
    return 0;
}
static int vgg_features_f32(const VggNet& net, const float* weights, const float* img, VggWsF& ws, Act<float> X0, Act<float>* A,
                            Act<float>* P, cudaStream_t s) {
    vgg_pack_input_f32_kernel<<<grid_for((int64_t)X0.N * X0.H * X0.W), NTHR, 0, s>>>(img, X0);
    MRISR_LAUNCH_CHECK();
    const Act<float> none{nullptr, 0, 0, 0, 0};
    Act<float> cur = X0;
    for (int i = 0; i < net.n_conv; ++i) {
        ConvArgs<float> a;
        a.src1 = cur; a.src2 = none; a.w = ws.wf[i]; a.bias = weights + net.b_off[i]; a.dst = A[i]; a.stat_part = nullptr; a.ks = 3;
        TRY(launch_conv_generic<float>(a, s));
        vgg_relu_mask_f32_kernel<<<grid_for((int64_t)A[i].N * A[i].H * A[i].W * (A[i].C / 8)), NTHR, 0, s>>>(A[i], none);
        MRISR_LAUNCH_CHECK();
        cur = A[i];
        if (net.pool_after[i]) {
            vgg_maxpool_kernel<float><<<grid_for((int64_t)P[i].N * P[i].H * P[i].W * (P[i].C / 8)), NTHR, 0, s>>>(A[i], P[i]);
            MRISR_LAUNCH_CHECK();
            cur = P[i];
        }
    }
    return 0;
}

}  // namespace
}  // namespace mrisr

using namespace mrisr;

extern "C" {

size_t mrisr_vgg_packed_bytes(const int* host_cfg, int ncfg) {
    VggNet net(host_cfg, ncfg);
    if (!net.ok) { set_error("vgg: bad config"); return 0; }
    return net.packed_bytes;
}

int64_t mrisr_vgg_weight_elems(const int* host_cfg, int ncfg) {
    VggNet net(host_cfg, ncfg);
    if (!net.ok) { set_error("vgg: bad config"); return -1; }
    return net.b_off[net.n_conv - 1] + net.cout[net.n_conv - 1];
}

int mrisr_vgg_pack_weights(const float* weights, const int* host_cfg, int ncfg, void* packed, size_t packed_bytes, void* stream) {
    VggNet net(host_cfg, ncfg);
    if (!net.ok || !weights || !packed) { set_error("vgg_pack_weights: bad argument"); return MRISR_E_BADARG; }
    if (packed_bytes < net.packed_bytes) { set_error("vgg_pack_weights: buffer too small (%zu < %zu)", packed_bytes, net.packed_bytes); return MRISR_E_WORKSPACE; }
    cudaStream_t s = (cudaStream_t)stream;
    for (int i = 0; i < net.n_conv; ++i) {
        bf16* wf = (bf16*)((char*)packed + net.wf_off[i]);
        bf16* wd = (bf16*)((char*)packed + net.wd_off[i]);
        if (i == 0) {
            vgg_pack_first_kernel<<<ceil_div(9 * net.cout[0] * kCin0, 256), 256, 0, s>>>(weights + net.w_off[0], wf, wd, net.cout[0]);
            MRISR_LAUNCH_CHECK();
        } else {
            TRY(launch_pack_weights_tc(weights + net.w_off[i], wf, wd, net.cout[i], net.cin[i], 3, s));
        }
    }
    return 0;
}

size_t mrisr_vgg_workspace_bytes(int N, int H, int W, const int* host_cfg, int ncfg, int grad) {
    VggNet net(host_cfg, ncfg);
    if (check_args(net, N, H, W)) return 0;
    return VggWs(nullptr, net, N, H, W, grad != 0).bytes;
}

int mrisr_vgg_features(const void* packed, const float* weights, const float* img, float* feat, void* workspace,
                       size_t workspace_bytes, int N, int H, int W, const int* host_cfg, int ncfg, void* stream) {
    VggNet net(host_cfg, ncfg);
    int rc = check_args(net, N, H, W);
    if (rc) return rc;
    if (!packed || !weights || !img || !feat || !workspace) { set_error("vgg_features: null pointer"); return MRISR_E_BADARG; }
    VggWs ws(workspace, net, N, H, W, false);
    if (ws.bytes > workspace_bytes) { set_error("vgg: workspace too small (%zu > %zu)", ws.bytes, workspace_bytes); return MRISR_E_WORKSPACE; }
    cudaStream_t s = (cudaStream_t)stream;
    TRY(vgg_features(net, packed, weights, img, ws.X0g, ws.Ag, ws.Pg, s));
    vgg_unpack_kernel<bf16><<<148 * 4, 256, 0, s>>>(ws.Ag[net.n_conv - 1], feat);
    MRISR_LAUNCH_CHECK();
    return 0;
}

int mrisr_vgg_perceptual_fwd_bwd(const void* packed, const float* weights, const float* pred, const float* target,
                                 float* dpred_accum, float* out1, void* workspace, size_t workspace_bytes, int N, int H,
                                 int W, const int* host_cfg, int ncfg, int loss_type, float grad_scale, void* stream) {
    VggNet net(host_cfg, ncfg);
    int rc = check_args(net, N, H, W);
    if (rc) return rc;
    if (!packed || !weights || !pred || !target || !out1 || !workspace) { set_error("vgg_perceptual: null pointer"); return MRISR_E_BADARG; }
    if (loss_type != 1 && loss_type != 2) { set_error("vgg_perceptual: loss_type must be 1 (L1) or 2 (MSE)"); return MRISR_E_BADARG; }
    const bool grad = dpred_accum != nullptr;
    VggWs ws(workspace, net, N, H, W, grad);
    if (ws.bytes > workspace_bytes) { set_error("vgg: workspace too small (%zu > %zu)", ws.bytes, workspace_bytes); return MRISR_E_WORKSPACE; }
    cudaStream_t s = (cudaStream_t)stream;
    const Act<bf16> none{nullptr, 0, 0, 0, 0};
    const int L = net.n_conv - 1;
    TRY(vgg_features(net, packed, weights, target, ws.X0t, ws.At, ws.Pt, s));      // no_grad branch, utils/losses.py:146-147
    TRY(vgg_features(net, packed, weights, pred, ws.X0g, ws.Ag, ws.Pg, s));
    {
        ProfScope prof_("vgg_feat_loss", s);
        const Act<bf16>& fg = ws.Ag[L];
        const int64_t numel = (int64_t)fg.N * fg.C * fg.H * fg.W;
        const int nb = grid_for(numel / 8);
        vgg_feat_loss_kernel<bf16><<<nb, NTHR, 0, s>>>(fg, ws.At[L], grad ? ws.dA[L] : none, ws.partial, loss_type, (float)(1.0 / (double)numel));
        MRISR_LAUNCH_CHECK();
        vgg_loss_finalize_kernel<<<1, 32, 0, s>>>(ws.partial, nb, 1.0 / (double)numel, out1);
        MRISR_LAUNCH_CHECK();
    }
    if (!grad) return 0;
    // dA[i] = gradient wrt the PRE-ReLU output of conv i
    for (int i = L; i >= 0; --i) {
        const bool pooled_input = i > 0 && net.pool_after[i - 1];
        ConvArgs<bf16> a;
        a.src1 = ws.dA[i]; a.src2 = none; a.w = (const bf16*)((const char*)packed + net.wd_off[i]);
        a.bias = nullptr; a.stat_part = nullptr; a.ks = 3; a.relu = false;
        if (i == 0) { a.dst = ws.dX0; a.mask = nullptr; }
        else if (pooled_input) { a.dst = ws.dP[i - 1]; a.mask = nullptr; }
        else { a.dst = ws.dA[i - 1]; a.mask = ws.Ag[i - 1].p; }                   // ReLU backward of conv i-1 in the epilogue
        TRY(launch_conv_tc(a, s));
        if (pooled_input) {
            ProfScope prof_("vgg_pool_bwd", s);
            const Act<bf16>& dp = ws.dP[i - 1];
            vgg_pool_bwd_kernel<bf16><<<grid_for((int64_t)dp.N * dp.H * dp.W * (dp.C / 8)), NTHR, 0, s>>>(ws.Ag[i - 1], dp, ws.dA[i - 1]);
            MRISR_LAUNCH_CHECK();
        }
    }
    {
        ProfScope prof_("vgg_dimg", s);
        vgg_dimg_kernel<bf16><<<grid_for((int64_t)N * H * W), NTHR, 0, s>>>(ws.dX0, dpred_accum, grad_scale);
        MRISR_LAUNCH_CHECK();
    }
    return 0;
}


/* ---- fp32 parity mode of the same term (CUDA cores; include/mrisr.h) ---- */
size_t mrisr_vgg_workspace_bytes_fp32(int N, int H, int W, const int* host_cfg, int ncfg, int grad) {
    VggNet net(host_cfg, ncfg);
    if (check_args_f32(net, N, H, W)) return 0;
    return VggWsF(nullptr, net, N, H, W, grad != 0).bytes;
}

int mrisr_vgg_features_fp32(const float* weights, const float* img, float* feat, void* workspace, size_t workspace_bytes, int N,
                            int H, int W, const int* host_cfg, int ncfg, void* stream) {
    VggNet net(host_cfg, ncfg);
    int rc = check_args_f32(net, N, H, W);
    if (rc) return rc;
    if (!weights || !img || !feat || !workspace) { set_error("vgg_features_fp32: null pointer"); return MRISR_E_BADARG; }
    VggWsF ws(workspace, net, N, H, W, false);
    if (ws.bytes > workspace_bytes) { set_error("vgg: workspace too small (%zu > %zu)", ws.bytes, workspace_bytes); return MRISR_E_WORKSPACE; }
    cudaStream_t s = (cudaStream_t)stream;
    TRY(vgg_pack_f32(net, weights, ws, s));
    TRY(vgg_features_f32(net, weights, img, ws, ws.X0g, ws.Ag, ws.Pg, s));
    vgg_unpack_kernel<float><<<148 * 4, 256, 0, s>>>(ws.Ag[net.n_conv - 1], feat);
    MRISR_LAUNCH_CHECK();
    return 0;
}

int mrisr_vgg_perceptual_fwd_bwd_fp32(const float* weights, const float* pred, const float* target, float* dpred_accum, float* out1,
                                      void* workspace, size_t workspace_bytes, int N, int H, int W, const int* host_cfg, int ncfg,
                                      int loss_type, float grad_scale, void* stream) {
    VggNet net(host_cfg, ncfg);
    int rc = check_args_f32(net, N, H, W);
    if (rc) return rc;
    if (!weights || !pred || !target || !out1 || !workspace) { set_error("vgg_perceptual_fp32: null pointer"); return MRISR_E_BADARG; }
    if (loss_type != 1 && loss_type != 2) { set_error("vgg_perceptual: loss_type must be 1 (L1) or 2 (MSE)"); return MRISR_E_BADARG; }
    const bool grad = dpred_accum != nullptr;
    VggWsF ws(workspace, net, N, H, W, grad);
    if (ws.bytes > workspace_bytes) { set_error("vgg: workspace too small (%zu > %zu)", ws.bytes, workspace_bytes); return MRISR_E_WORKSPACE; }
    cudaStream_t s = (cudaStream_t)stream;
    const Act<float> none{nullptr, 0, 0, 0, 0};
    const int L = net.n_conv - 1;
    TRY(vgg_pack_f32(net, weights, ws, s));
    TRY(vgg_features_f32(net, weights, target, ws, ws.X0t, ws.At, ws.Pt, s));
    TRY(vgg_features_f32(net, weights, pred, ws, ws.X0g, ws.Ag, ws.Pg, s));
    {
        const Act<float>& fg = ws.Ag[L];
        const int64_t numel = (int64_t)fg.N * fg.C * fg.H * fg.W;
        const int nb = grid_for(numel / 8);
        vgg_feat_loss_kernel<float><<<nb, NTHR, 0, s>>>(fg, ws.At[L], grad ? ws.dA[L] : none, ws.partial, loss_type, (float)(1.0 / (double)numel));
        MRISR_LAUNCH_CHECK();
        vgg_loss_finalize_kernel<<<1, 32, 0, s>>>(ws.partial, nb, 1.0 / (double)numel, out1);
        MRISR_LAUNCH_CHECK();
    }
    if (!grad) return 0;
    for (int i = L; i >= 0; --i) {
        const bool pooled_input = i > 0 && net.pool_after[i - 1];
        ConvArgs<float> a;
        a.src1 = ws.dA[i]; a.src2 = none; a.w = ws.wd[i]; a.bias = nullptr; a.stat_part = nullptr; a.ks = 3;
        a.dst = i == 0 ? ws.dX0 : (pooled_input ? ws.dP[i - 1] : ws.dA[i - 1]);
        TRY(launch_conv_generic<float>(a, s));
        if (i > 0 && !pooled_input) {
            const Act<float>& d = ws.dA[i - 1];
            vgg_relu_mask_f32_kernel<<<grid_for((int64_t)d.N * d.H * d.W * (d.C / 8)), NTHR, 0, s>>>(d, ws.Ag[i - 1]);
            MRISR_LAUNCH_CHECK();
        } else if (pooled_input) {
            const Act<float>& dp = ws.dP[i - 1];
            vgg_pool_bwd_kernel<float><<<grid_for((int64_t)dp.N * dp.H * dp.W * (dp.C / 8)), NTHR, 0, s>>>(ws.Ag[i - 1], dp, ws.dA[i - 1]);
            MRISR_LAUNCH_CHECK();
        }
    }
    vgg_dimg_kernel<float><<<grid_for((int64_t)N * H * W), NTHR, 0, s>>>(ws.dX0, dpred_accum, grad_scale);
    MRISR_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
