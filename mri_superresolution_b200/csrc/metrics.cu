// Batched evaluation side of the path:
//   scripts/test_comparison.py:164-202   calculate_metrics: SSIM (loss.cu), MSE, RMSE, MAE, PSNR per image
//   scripts/test_comparison.py:92-134    upscale_with_interpolation: cv2.resize x2 INTER_LINEAR / INTER_CUBIC on uint8, the
//                                        "sharp bilinear" 3x3 filter2D, result / 255
//   scripts/infer.py:285-314             skimage.exposure.match_histograms(output, reference) (CDF matching)
// The cv2 paths are fixed-point integer arithmetic and are reproduced bit for bit against cv2 4.13 (oracle/eval_oracle.py states
// the arithmetic; tests/golden/eval.npz holds cv2's own outputs).  One thread per output pixel; everything here is a small
// HBM-bound pass over at most a few hundred MB.
#include "kernels.cuh"
#include "../../include/mrisr.h"
#include <algorithm>
#include <cub/device/device_radix_sort.cuh>

namespace mrisr {
namespace {
inline int grid1d(long long total, int thr) {
    long long b = (total + thr - 1) / thr;
    if (b > 148 * 32) b = 148 * 32;
    return b < 1 ? 1 : (int)b;
}

// ---- metrics: per-image sum of squared and absolute differences in double ------------------------------------------------
constexpr int MT = 512;
__global__ void __launch_bounds__(MT) diff_sums_kernel(const float* __restrict__ a, const float* __restrict__ b, double* __restrict__ part,
                                                       long long hw, int blocks_per_image) {
    __shared__ double s_w[2][MT / 32];
    const int n = blockIdx.x / blocks_per_image, blk = blockIdx.x % blocks_per_image;
    const float* pa = a + (long long)n * hw;
    const float* pb = b + (long long)n * hw;
    double s2 = 0.0, s1 = 0.0;
    for (long long i = (long long)blk * MT + threadIdx.x; i < hw; i += (long long)blocks_per_image * MT) {
        const float d = __fsub_rn(pa[i], pb[i]);                   // the reference subtracts in float32
        s2 += (double)__fmul_rn(d, d);
        s1 += (double)fabsf(d);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { s2 += __shfl_xor_sync(0xffffffffu, s2, o); s1 += __shfl_xor_sync(0xffffffffu, s1, o); }
    if ((threadIdx.x & 31) == 0) { s_w[0][threadIdx.x >> 5] = s2; s_w[1][threadIdx.x >> 5] = s1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        s2 = s1 = 0.0;
        for (int w = 0; w < MT / 32; ++w) { s2 += s_w[0][w]; s1 += s_w[1][w]; }
        part[2 * (long long)blockIdx.x] = s2;
        part[2 * (long long)blockIdx.x + 1] = s1;
    }
}
// out[n] = (mse, rmse, mae, psnr): psnr = 100 when mse < 1e-10 (test_comparison.py:191-192), else 10 log10(1 / mse)
__global__ void metrics_finalize_kernel(const double* __restrict__ part, int blocks_per_image, double inv_hw, float* __restrict__ out) {
    const int n = blockIdx.x;
    double s2 = 0.0, s1 = 0.0;
    for (int i = threadIdx.x; i < blocks_per_image; i += 32) { s2 += part[2 * ((long long)n * blocks_per_image + i)]; s1 += part[2 * ((long long)n * blocks_per_image + i) + 1]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { s2 += __shfl_xor_sync(0xffffffffu, s2, o); s1 += __shfl_xor_sync(0xffffffffu, s1, o); }
    if (threadIdx.x == 0) {
        const double mse = s2 * inv_hw, mae = s1 * inv_hw;
        out[4 * n] = (float)mse;
        out[4 * n + 1] = (float)sqrt(mse);
        out[4 * n + 2] = (float)mae;
        out[4 * n + 3] = mse < 1e-10 ? 100.f : (float)(10.0 * log10(1.0 / mse));
    }
}

// ---- cv2.resize x2 on uint8 ------------------------------------------------------------------------------------------
// Source coordinate of destination index d at scale 1/2: f = (d + 0.5) * 0.5 - 0.5 -> s = floor(f), fraction 0.25 / 0.75.
__device__ __forceinline__ void src_coord(int d, int& s, int& frac4) {          // frac4 = 4 * fraction (1 or 3)
    s = (d + 1) / 2 - 1;                                                         // floor((2d - 1) / 4)
    frac4 = (d & 1) ? 1 : 3;
}
// INTER_LINEAR (resize.cpp: HResizeLinear + VResizeLinear<uchar, int, short>): coefficients * 2048 as shorts; horizontally the
// border clamps the INDEX and zeroes the fraction, vertically only the row index is clamped (both rows then coincide);
// dst = (((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2.
__device__ __forceinline__ int hlinear(const uint8_t* row, int W, int dx) {
    int sx, f4;
    src_coord(dx, sx, f4);
    int c1 = 512 * f4;
    if (sx < 0) { sx = 0; c1 = 0; }
    if (sx >= W - 1) { sx = W - 1; c1 = 0; }
    const int x1 = sx + 1 < W ? sx + 1 : W - 1;
    return (int)row[sx] * (2048 - c1) + (int)row[x1] * c1;
}
// INTER_CUBIC (A = -0.75): taps * 2048 at fraction 0.25 = (-216, 1800, 536, -72), at 0.75 mirrored; indices clamped.
__device__ __forceinline__ void cubic_taps(int f4, int (&c)[4]) {
    if (f4 == 1) { c[0] = -216; c[1] = 1800; c[2] = 536; c[3] = -72; }
    else { c[0] = -72; c[1] = 536; c[2] = 1800; c[3] = -216; }
}
__device__ __forceinline__ int hcubic(const uint8_t* row, int W, int dx) {
    int sx, f4, c[4];
    src_coord(dx, sx, f4);
    cubic_taps(f4, c);
    int acc = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int xi = sx - 1 + k;
        xi = xi < 0 ? 0 : (xi > W - 1 ? W - 1 : xi);
        acc += (int)row[xi] * c[k];
    }
    return acc;
}

template <bool CUBIC>
__global__ void __launch_bounds__(256) upscale2x_u8_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int H, int W, long long total) {
    const int W2 = 2 * W, H2 = 2 * H;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int dx = (int)(i % W2);
        const long long r = i / W2;
        const int dy = (int)(r % H2);
        const uint8_t* img = src + (r / H2) * (long long)H * W;
        int sy, f4;
        src_coord(dy, sy, f4);
        int v;
        if (!CUBIC) {
            const int y0 = sy < 0 ? 0 : (sy > H - 1 ? H - 1 : sy), y1 = sy + 1 < 0 ? 0 : (sy + 1 > H - 1 ? H - 1 : sy + 1);
            const int b1 = 512 * f4, b0 = 2048 - b1;
            const int S0 = hlinear(img + (long long)y0 * W, W, dx), S1 = hlinear(img + (long long)y1 * W, W, dx);
            v = (((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2;
        } else {
            int c[4], S[4];
            cubic_taps(f4, c);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int yi = sy - 1 + k;
                yi = yi < 0 ? 0 : (yi > H - 1 ? H - 1 : yi);
                S[k] = hcubic(img + (long long)yi * W, W, dx);
            }
            // cv2's x86 build: the vector body of VResizeCubic (groups of 4 columns) evaluates in float32,
            // s0*b0 + (s1*b1 + (s2*b2 + s3*b3)) with b = tap / 2^22, rounded half-to-even; the scalar tail of a row uses the
            // fixed-point cast (sum + 2^21) >> 22.  They differ only on exact .5 ties and float32 rounding of the products.
            if (dx < (W2 / 4) * 4) {
                const float sc = 1.f / 4194304.f;
                float rr = __fmul_rn((float)S[3], __fmul_rn((float)c[3], sc));
                rr = __fadd_rn(__fmul_rn((float)S[2], __fmul_rn((float)c[2], sc)), rr);
                rr = __fadd_rn(__fmul_rn((float)S[1], __fmul_rn((float)c[1], sc)), rr);
                rr = __fadd_rn(__fmul_rn((float)S[0], __fmul_rn((float)c[0], sc)), rr);
                v = __float2int_rn(rr);
            } else {
                long long acc = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k) acc += (long long)S[k] * c[k];
                v = (int)((acc + (1ll << 21)) >> 22);
            }
        }
        dst[i] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
    }
}

// optional filter2D([[-1,-1,-1],[-1,9,-1],[-1,-1,-1]], BORDER_REFLECT_101, saturate to uint8), then / 255
__global__ void __launch_bounds__(256) sharpen_to_float_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst, uint8_t* __restrict__ dst_u8,
                                                               int H, int W, long long total, int sharpen) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int x = (int)(i % W);
        const long long r = i / W;
        const int y = (int)(r % H);
        const uint8_t* img = src + (r / H) * (long long)H * W;
        int v = img[(long long)y * W + x];
        if (sharpen) {
            int acc = 10 * v;
#pragma unroll
            for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
                for (int dx = -1; dx <= 1; ++dx) {
                    int yy = y + dy, xx = x + dx;
                    yy = yy < 0 ? -yy : (yy > H - 1 ? 2 * (H - 1) - yy : yy);
                    xx = xx < 0 ? -xx : (xx > W - 1 ? 2 * (W - 1) - xx : xx);
                    if (H == 1) yy = 0;
                    if (W == 1) xx = 0;
                    acc -= img[(long long)yy * W + xx];
                }
            v = acc < 0 ? 0 : (acc > 255 ? 255 : acc);
        }
        if (dst_u8) dst_u8[i] = (uint8_t)v;
        if (dst) dst[i] = __fdiv_rn((float)v, 255.f);
    }
}

// ---- skimage.exposure.match_histograms (scripts/infer.py:303-307), one (source, reference) pair per image -----------------
//   src_values, src_idx, src_counts = unique(source);  tmpl_values, tmpl_counts = unique(template)
//   src_q = cumsum(src_counts) / source.size;  tmpl_q = cumsum(tmpl_counts) / template.size
//   out = interp(src_q, tmpl_q, tmpl_values)[src_idx]                      (float64 arithmetic)
// Both images are sorted once (cub::DeviceRadixSort on 64-bit keys = image index : order-preserving float bits; a library sort
// for a post-processing step outside the timed path); the cumulative count of a value is then the end of its run.
__device__ __forceinline__ unsigned f2ord(float f) {
    const unsigned b = __float_as_uint(f == 0.f ? 0.f : f);                       // -0 == +0 for np.unique
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned o) { return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o); }

__global__ void __launch_bounds__(256) hm_keys_kernel(const float* __restrict__ img, unsigned long long* __restrict__ keys,
                                                      unsigned* __restrict__ idx, long long hw, long long total) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        keys[i] = ((unsigned long long)(i / hw) << 32) | f2ord(img[i]);
        if (idx) idx[i] = (unsigned)(i % hw);
    }
}
// last position (within the image's sorted block [0, m)) holding the same key as position p
__device__ __forceinline__ long long run_end(const unsigned long long* __restrict__ k, long long m, long long p) {
    const unsigned long long key = k[p];
    long long lo = p, hi = m - 1;                       // invariant: k[lo] == key
    while (lo < hi) {
        const long long mid = (lo + hi + 1) >> 1;
        if (k[mid] == key) lo = mid; else hi = mid - 1;
    }
    return lo;
}
__device__ __forceinline__ long long run_begin(const unsigned long long* __restrict__ k, long long p) {
    const unsigned long long key = k[p];
    long long lo = 0, hi = p;
    while (lo < hi) {
        const long long mid = (lo + hi) >> 1;
        if (k[mid] == key) hi = mid; else lo = mid + 1;
    }
    return lo;
}
__global__ void __launch_bounds__(256) hm_apply_kernel(const unsigned long long* __restrict__ skeys, const unsigned* __restrict__ sidx,
                                                       const unsigned long long* __restrict__ tkeys, float* __restrict__ out,
                                                       long long hw, long long thw, long long total) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long n = i / hw, p = i % hw;
        const unsigned long long* sk = skeys + n * hw;
        const unsigned long long* tk = tkeys + n * thw;
        const double q = (double)(run_end(sk, hw, p) + 1) / (double)hw;           // cumsum(src_counts)[u] / source.size
        // largest template position J with (J + 1) / thw <= q
        long long lo = -1, hi = thw - 1;
        while (lo < hi) {
            const long long mid = (lo + hi + 1) >> 1;
            if ((double)(mid + 1) / (double)thw <= q) lo = mid; else hi = mid - 1;
        }
        double r;
        if (lo < 0) {
            r = (double)ord2f((unsigned)tk[0]);                                   // below xp[0]: fp[0]
        } else {
            long long e = run_end(tk, thw, lo);
            if (e != lo) e = run_begin(tk, lo) - 1;                                // the run containing J is not complete at J
            if (e < 0) {
                r = (double)ord2f((unsigned)tk[0]);
            } else if (e == thw - 1) {
                r = (double)ord2f((unsigned)tk[e]);                                // at / beyond xp[-1]
            } else {
                const double x0 = (double)(e + 1) / (double)thw, f0 = (double)ord2f((unsigned)tk[e]);
                const long long e1 = run_end(tk, thw, e + 1);
                const double x1 = (double)(e1 + 1) / (double)thw, f1 = (double)ord2f((unsigned)tk[e + 1]);
                if (x0 == q) r = f0;
                else {
                    const double slope = (f1 - f0) / (x1 - x0);
                    r = slope * (q - x0) + f0;
                }
            }
        }
        out[n * hw + sidx[i]] = (float)r;
    }
}

struct HmLayout { size_t skeys_a, skeys_b, sidx_a, sidx_b, tkeys_a, tkeys_b, cub, total; };
inline size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }
HmLayout hm_layout(long long ns, long long nt) {
    HmLayout L;
    size_t o = 0;
    L.skeys_a = o; o += al256((size_t)ns * 8);
    L.skeys_b = o; o += al256((size_t)ns * 8);
    L.sidx_a = o; o += al256((size_t)ns * 4);
    L.sidx_b = o; o += al256((size_t)ns * 4);
    L.tkeys_a = o; o += al256((size_t)nt * 8);
    L.tkeys_b = o; o += al256((size_t)nt * 8);
    size_t t1 = 0, t2 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, t1, (const unsigned long long*)nullptr, (unsigned long long*)nullptr, (const unsigned*)nullptr,
                                    (unsigned*)nullptr, (int)ns);
    cub::DeviceRadixSort::SortKeys(nullptr, t2, (const unsigned long long*)nullptr, (unsigned long long*)nullptr, (int)nt);
    L.cub = o; o += al256(t1 > t2 ? t1 : t2);
    L.total = o;
    return L;
}
}  // namespace
}  // namespace mrisr

using namespace mrisr;

extern "C" {

size_t mrisr_image_metrics_scratch_bytes(int N, int H, int W) {
    if (N <= 0 || H <= 0 || W <= 0) return 0;
    const int bpi = (int)std::min<long long>(64, ((long long)H * W + MT * 8 - 1) / (MT * 8));
    return (size_t)N * (bpi < 1 ? 1 : bpi) * 2 * sizeof(double);
}

int mrisr_image_metrics(const float* reference, const float* image, float* out4, void* scratch, int N, int H, int W, void* stream) {
    if (!reference || !image || !out4 || !scratch || N <= 0 || H <= 0 || W <= 0) { set_error("bad argument"); return MRISR_E_BADARG; }
    cudaStream_t s = (cudaStream_t)stream;
    ProfScope prof_("image_metrics", s);
    const long long hw = (long long)H * W;
    int bpi = (int)std::min<long long>(64, (hw + MT * 8 - 1) / (MT * 8));
    if (bpi < 1) bpi = 1;
    diff_sums_kernel<<<N * bpi, MT, 0, s>>>(reference, image, (double*)scratch, hw, bpi);
    MRISR_LAUNCH_CHECK();
    metrics_finalize_kernel<<<N, 32, 0, s>>>((const double*)scratch, bpi, 1.0 / (double)hw, out4);
    MRISR_LAUNCH_CHECK();
    return 0;
}

int mrisr_upscale2x_u8(const uint8_t* src, float* dst, uint8_t* dst_u8, uint8_t* tmp, int N, int H, int W, int method, void* stream) {
    if (!src || (!dst && !dst_u8) || !tmp || N <= 0 || H <= 0 || W <= 0) { set_error("bad argument"); return MRISR_E_BADARG; }
    if (method < 0 || method > 2) { set_error("upscale2x: method must be 0 (bilinear), 1 (sharp_bilinear) or 2 (bicubic)"); return MRISR_E_BADARG; }
    cudaStream_t s = (cudaStream_t)stream;
    ProfScope prof_("upscale2x_u8", s);
    const long long total = (long long)N * 4 * H * W;
    if (method == 2) upscale2x_u8_kernel<true><<<grid1d(total, 256), 256, 0, s>>>(src, tmp, H, W, total);
    else upscale2x_u8_kernel<false><<<grid1d(total, 256), 256, 0, s>>>(src, tmp, H, W, total);
    MRISR_LAUNCH_CHECK();
    sharpen_to_float_kernel<<<grid1d(total, 256), 256, 0, s>>>(tmp, dst, dst_u8, 2 * H, 2 * W, total, method == 1 ? 1 : 0);
    MRISR_LAUNCH_CHECK();
    return 0;
}

size_t mrisr_match_histograms_scratch_bytes(int N, int H, int W, int Href, int Wref) {
    if (N <= 0 || H <= 0 || W <= 0 || Href <= 0 || Wref <= 0) return 0;
    const long long ns = (long long)N * H * W, nt = (long long)N * Href * Wref;
    if (ns >= (1ll << 31) || nt >= (1ll << 31)) return 0;
    return hm_layout(ns, nt).total;
}

int mrisr_match_histograms(const float* source, const float* reference, float* out, void* scratch, size_t scratch_bytes, int N, int H,
                           int W, int Href, int Wref, void* stream) {
    if (!source || !reference || !out || !scratch || N <= 0 || H <= 0 || W <= 0 || Href <= 0 || Wref <= 0) { set_error("bad argument"); return MRISR_E_BADARG; }
    const long long hw = (long long)H * W, thw = (long long)Href * Wref, ns = hw * N, nt = thw * N;
    if (ns >= (1ll << 31) || nt >= (1ll << 31)) { set_error("match_histograms: at most 2^31 - 1 pixels per call"); return MRISR_E_SHAPE; }
    const HmLayout L = hm_layout(ns, nt);
    if (scratch_bytes < L.total) { set_error("match_histograms: scratch too small: need %zu bytes, got %zu", L.total, scratch_bytes); return MRISR_E_WORKSPACE; }
    cudaStream_t s = (cudaStream_t)stream;
    ProfScope prof_("match_histograms", s);
    char* b = (char*)scratch;
    auto* ska = (unsigned long long*)(b + L.skeys_a); auto* skb = (unsigned long long*)(b + L.skeys_b);
    auto* sia = (unsigned*)(b + L.sidx_a); auto* sib = (unsigned*)(b + L.sidx_b);
    auto* tka = (unsigned long long*)(b + L.tkeys_a); auto* tkb = (unsigned long long*)(b + L.tkeys_b);
    size_t cub_bytes = L.total - L.cub;
    hm_keys_kernel<<<grid1d(ns, 256), 256, 0, s>>>(source, ska, sia, hw, ns);
    MRISR_LAUNCH_CHECK();
    hm_keys_kernel<<<grid1d(nt, 256), 256, 0, s>>>(reference, tka, nullptr, thw, nt);
    MRISR_LAUNCH_CHECK();
    MRISR_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(b + L.cub, cub_bytes, ska, skb, sia, sib, (int)ns, 0, 64, s));
    cub_bytes = L.total - L.cub;
    MRISR_CUDA_CHECK(cub::DeviceRadixSort::SortKeys(b + L.cub, cub_bytes, tka, tkb, (int)nt, 0, 64, s));
    hm_apply_kernel<<<grid1d(ns, 256), 256, 0, s>>>(skb, sib, tkb, out, hw, thw, ns);
    MRISR_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
