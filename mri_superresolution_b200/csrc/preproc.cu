// On-device input pipeline: what the reference does per image on the host with PIL / numpy before a
// slice reaches the model, restated as batched kernels over dense uint8 [N][H][W] slices.
//   scripts/infer.py:97-130      preprocess_image: np.percentile(0.5 / 99.5) clip + min-max to [0,1]
//   utils/dataset.py:124-125     transforms.ToTensor(): uint8 / 255 -> float32
//   utils/dataset.py:138-175     augment_pair: hflip, rotate (PIL NEAREST, fill = mean intensity), brightness, contrast
//                                (PIL ImageEnhance = Image.blend with black / the mean-gray image), Gaussian noise on the LR image
// All of it is byte / index arithmetic reproduced bit for bit (tests/test_gpu_preproc.py against goldens made by the
// reference's own code): the percentile is numpy's 'linear' method evaluated in float32 exactly as np.percentile does for a
// float32 array (a uint8 slice has 256 levels, so the two order statistics come from a 256-bin histogram: no sort); the
// rotation is Pillow's 16.16 fixed-point affine_fixed loop; the blends truncate like Pillow's Blend.c.
// HBM-bound byte work: one CTA per slice for the reductions, grid-stride elementwise passes otherwise.
#include "kernels.cuh"
#include "../../include/mrisr.h"

namespace mrisr {
namespace {
constexpr int PT = 1024;       // threads of the per-slice kernels

__device__ __forceinline__ unsigned long long block_sum_u64(unsigned long long v, unsigned long long* s_w) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = v;
    __syncthreads();
    unsigned long long t = 0;
    if (threadIdx.x == 0)
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_w[w];
    return t;                   // valid in thread 0
}

// ---- scripts/infer.py:106-116 ---------------------------------------------------------------------------------------
// np.percentile(float32 array, q): q32 = float32(q) / float32(100); virtual index = float32(n - 1) * q32 (float32);
// lo = floor, gamma = virt - lo (float32); result = a + (b - a) * gamma, or b - (b - a) * (1 - gamma) when gamma >= 0.5, all
// in float32 (numpy lib/_function_base_impl.py: _quantile, _lerp).  a, b = the order statistics lo and lo + 1.
__device__ __forceinline__ float np_percentile_f32(const unsigned* cum, long long n, float q) {
    const float q32 = __fdiv_rn(q, 100.f);
    const float virt = __fmul_rn((float)(n - 1), q32);
    const float fl = floorf(virt);
    long long lo = (long long)fl;
    if (lo > n - 1) lo = n - 1;
    const long long hi = lo + 1 < n ? lo + 1 : n - 1;
    const float g = __fsub_rn(virt, fl);
    int ia = 0, ib = 0;
    for (int v = 0; v < 256; ++v) {                      // first value whose cumulative count exceeds the index
        if (cum[v] <= (unsigned long long)lo) ia = v + 1;
        if (cum[v] <= (unsigned long long)hi) ib = v + 1;
    }
    const float a = (float)ia, b = (float)ib, d = __fsub_rn(b, a);
    float r = __fadd_rn(a, __fmul_rn(d, g));
    if (g >= 0.5f) r = __fsub_rn(b, __fmul_rn(d, __fsub_rn(1.f, g)));
    return r;
}

__global__ void __launch_bounds__(PT) percentile_minmax_kernel(const uint8_t* __restrict__ src, float* __restrict__ minmax,
                                                               long long hw, float q_lo, float q_hi) {
    __shared__ unsigned hist[256];
    __shared__ unsigned cum[256];
    const uint8_t* p = src + (long long)blockIdx.x * hw;
    if (threadIdx.x < 256) hist[threadIdx.x] = 0;
    __syncthreads();
    // 16-byte loads where the slice base is aligned (hw % 16 == 0 for every shape the path sees), byte tail otherwise
    const bool vec = (hw % 16 == 0) && ((reinterpret_cast<uintptr_t>(p) & 15) == 0);
    if (vec) {
        const uint4* p4 = reinterpret_cast<const uint4*>(p);
        for (long long i = threadIdx.x; i < hw / 16; i += PT) {
            const uint4 v = p4[i];
            const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int b = 0; b < 4; ++b) atomicAdd(&hist[(w[k] >> (8 * b)) & 255u], 1u);
        }
    } else {
        for (long long i = threadIdx.x; i < hw; i += PT) atomicAdd(&hist[p[i]], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned c = 0;
        for (int v = 0; v < 256; ++v) { c += hist[v]; cum[v] = c; }
        minmax[2 * blockIdx.x] = np_percentile_f32(cum, hw, q_lo);
        minmax[2 * blockIdx.x + 1] = np_percentile_f32(cum, hw, q_hi);
    }
}

// clip to [min, max] then (x - min) / (max - min) in float32 when max > min; else the clipped raw values (infer.py:115-116)
__global__ void __launch_bounds__(256) percentile_apply_kernel(const uint8_t* __restrict__ src, const float* __restrict__ minmax,
                                                               float* __restrict__ dst, long long hw, long long total) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long n = i / hw;
        const float lo = minmax[2 * n], hi = minmax[2 * n + 1];
        float v = (float)src[i];
        v = fminf(fmaxf(v, lo), hi);
        if (hi > lo) v = __fdiv_rn(__fsub_rn(v, lo), __fsub_rn(hi, lo));
        dst[i] = v;
    }
}

__global__ void __launch_bounds__(256) u8_to_float_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst, long long total) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x)
        dst[i] = __fdiv_rn((float)src[i], 255.f);                    // torchvision to_tensor: .to(float32).div(255)
}

// ---- utils/dataset.py:138-175 ---------------------------------------------------------------------------------------
// params: MRISR_AUG_PARAM_INTS int32 per sample (include/mrisr.h)
struct AugParams {
    int flip, rotate, a0, a1, a2, a3, a4, a5, brightness, brightness_bits, contrast, contrast_bits, noise, noise_std_bits, seed_lo, seed_hi;
};
static_assert(sizeof(AugParams) == 4 * MRISR_AUG_PARAM_INTS, "parameter block layout");

// stage 1: per-slice sum of the source (rotation fill = int(mean(x / 255) * 255), dataset.py:150-151); zeroes the stage-2 sum
__global__ void __launch_bounds__(PT) aug_sum_kernel(const uint8_t* __restrict__ src, unsigned long long* __restrict__ sums, long long hw) {
    __shared__ unsigned long long s_w[PT / 32];
    const uint8_t* p = src + (long long)blockIdx.x * hw;
    unsigned long long acc = 0;
    for (long long i = threadIdx.x; i < hw; i += PT) acc += p[i];
    const unsigned long long t = block_sum_u64(acc, s_w);
    if (threadIdx.x == 0) { sums[2 * blockIdx.x] = t; sums[2 * blockIdx.x + 1] = 0; }
}

// Pillow Blend.c: out = in1 + alpha * (in2 - in1) in float, truncated; extrapolation (alpha outside [0,1]) clips first
__device__ __forceinline__ int pil_blend(int degenerate, int v, float alpha) {
    const float t = __fadd_rn((float)degenerate, __fmul_rn(alpha, (float)(v - degenerate)));
    if (t <= 0.f) return 0;
    if (t >= 255.f) return 255;
    return (int)t;
}

// stage 2: hflip -> rotate (Pillow Geometry.c affine_fixed, NEAREST) -> brightness; uint8 out + per-slice sum of the result
__global__ void __launch_bounds__(256) aug_gather_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ tmp,
                                                         const AugParams* __restrict__ params, unsigned long long* __restrict__ sums,
                                                         int H, int W, int blocks_per_slice) {
    __shared__ unsigned long long s_w[8];
    const int n = blockIdx.x / blocks_per_slice, b = blockIdx.x % blocks_per_slice;
    const AugParams pr = params[n];
    const long long hw = (long long)H * W;
    const uint8_t* p = src + n * hw;
    uint8_t* q = tmp + n * hw;
    int fill = 0;
    if (pr.rotate) {
        const float m = (float)((double)sums[2 * n] / (255.0 * (double)hw));       // float32 mean of x / 255
        fill = (int)((double)m * 255.0);
    }
    const float bf = __int_as_float(pr.brightness_bits);
    unsigned long long acc = 0;
    for (long long i = (long long)b * 256 + threadIdx.x; i < hw; i += (long long)blocks_per_slice * 256) {
        const int y = (int)(i / W), x = (int)(i % W);
        int v;
        if (pr.rotate) {
            const long long xx = (long long)pr.a2 + (long long)pr.a1 * y + (long long)pr.a0 * x;
            const long long yy = (long long)pr.a5 + (long long)pr.a4 * y + (long long)pr.a3 * x;
            const long long xin = xx >> 16, yin = yy >> 16;
            if (xin >= 0 && xin < W && yin >= 0 && yin < H) v = p[yin * W + (pr.flip ? W - 1 - xin : xin)];
            else v = fill;
        } else {
            v = p[(long long)y * W + (pr.flip ? W - 1 - x : x)];
        }
        if (pr.brightness) v = pil_blend(0, v, bf);
        q[i] = (uint8_t)v;
        acc += (unsigned)v;
    }
    const unsigned long long t = block_sum_u64(acc, s_w);
    if (threadIdx.x == 0) atomicAdd(&sums[2 * n + 1], t);
}

// Philox-4x32-10 (counter = pixel index, key = per-sample seed) + Box-Muller: the device-side noise source when no explicit
// noise array is given (the reference draws from numpy's global Mersenne-Twister stream, which no GPU generator reproduces)
__device__ __forceinline__ void philox4x32_10(unsigned c0, unsigned c1, unsigned c2, unsigned c3, unsigned k0, unsigned k1, unsigned (&out)[4]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const unsigned hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const unsigned hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
__device__ __forceinline__ float philox_normal(long long i, unsigned k0, unsigned k1) {
    unsigned r[4];
    philox4x32_10((unsigned)i, (unsigned)((unsigned long long)i >> 32), 0u, 0u, k0, k1, r);
    const float u1 = ((float)(r[0] >> 8) + 0.5f) * (1.f / 16777216.f), u2 = ((float)(r[1] >> 8) + 0.5f) * (1.f / 16777216.f);
    return sqrtf(-2.f * __logf(u1)) * __cosf(6.28318530717958647692f * u2);
}

// stage 3: contrast (mean of the stage-2 image: int(sum / count + 0.5)) -> noise (clip(x + noise, 0, 255) truncated, :168-170)
// -> ToTensor
__global__ void __launch_bounds__(256) aug_finish_kernel(const uint8_t* __restrict__ tmp, float* __restrict__ dst, uint8_t* __restrict__ dst_u8,
                                                         const AugParams* __restrict__ params, const unsigned long long* __restrict__ sums,
                                                         const double* __restrict__ noise, long long hw, long long total) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long n = i / hw;
        const AugParams pr = params[n];
        int v = tmp[i];
        if (pr.contrast) {
            const int mean = (int)((double)sums[2 * n + 1] / (double)hw + 0.5);
            v = pil_blend(mean, v, __int_as_float(pr.contrast_bits));
        }
        if (pr.noise) {
            const double sd = (double)__int_as_float(pr.noise_std_bits);
            const double e = noise ? noise[i] : sd * (double)philox_normal(i - n * hw, (unsigned)pr.seed_lo, (unsigned)pr.seed_hi);
            double t = (double)v + e;
            t = t < 0.0 ? 0.0 : (t > 255.0 ? 255.0 : t);
            v = (int)t;
        }
        if (dst_u8) dst_u8[i] = (uint8_t)v;
        if (dst) dst[i] = __fdiv_rn((float)v, 255.f);
    }
}

inline int grid1d(long long total, int thr) {
    long long b = (total + thr - 1) / thr;
    if (b > 148 * 32) b = 148 * 32;
    return b < 1 ? 1 : (int)b;
}
}  // namespace
}  // namespace mrisr

using namespace mrisr;

extern "C" {

int mrisr_u8_to_float(const uint8_t* src, float* dst, int64_t count, void* stream) {
    if (!src || !dst || count <= 0) { set_error("bad argument"); return MRISR_E_BADARG; }
    ProfScope prof_("u8_to_float", (cudaStream_t)stream);
    u8_to_float_kernel<<<grid1d(count, 256), 256, 0, (cudaStream_t)stream>>>(src, dst, count);
    MRISR_LAUNCH_CHECK();
    return 0;
}

int mrisr_percentile_normalize_u8(const uint8_t* src, float* dst, float* minmax, int N, int H, int W, float lower_percentile,
                                  float upper_percentile, void* stream) {
    if (!src || !dst || !minmax || N <= 0 || H <= 0 || W <= 0) { set_error("bad argument"); return MRISR_E_BADARG; }
    if (!(lower_percentile >= 0.f && upper_percentile <= 100.f && lower_percentile <= upper_percentile)) {
        set_error("percentiles must satisfy 0 <= lower <= upper <= 100"); return MRISR_E_BADARG;
    }
    if ((long long)H * W >= (1ll << 24)) { set_error("slices of 2^24 pixels or more are not supported (float32 order-statistic index)"); return MRISR_E_SHAPE; }
    cudaStream_t s = (cudaStream_t)stream;
    ProfScope prof_("percentile_normalize", s);
    const long long hw = (long long)H * W;
    percentile_minmax_kernel<<<N, PT, 0, s>>>(src, minmax, hw, lower_percentile, upper_percentile);
    MRISR_LAUNCH_CHECK();
    percentile_apply_kernel<<<grid1d(hw * N, 256), 256, 0, s>>>(src, minmax, dst, hw, hw * N);
    MRISR_LAUNCH_CHECK();
    return 0;
}

int mrisr_augment_u8(const uint8_t* src, float* dst, uint8_t* dst_u8, uint8_t* tmp, uint64_t* sums, const int32_t* params,
                     const double* noise, int N, int H, int W, void* stream) {
    if (!src || (!dst && !dst_u8) || !tmp || !sums || !params || N <= 0 || H <= 0 || W <= 0) { set_error("bad argument"); return MRISR_E_BADARG; }
    if (H >= 0x4000 || W >= 0x4000) { set_error("augment: Pillow's fixed-point affine path covers images below 16384 pixels per side"); return MRISR_E_SHAPE; }
    cudaStream_t s = (cudaStream_t)stream;
    ProfScope prof_("augment", s);
    const long long hw = (long long)H * W;
    const AugParams* pr = reinterpret_cast<const AugParams*>(params);
    unsigned long long* sm = reinterpret_cast<unsigned long long*>(sums);
    aug_sum_kernel<<<N, PT, 0, s>>>(src, sm, hw);
    MRISR_LAUNCH_CHECK();
    int bps = (int)((hw + 256 * 16 - 1) / (256 * 16));
    if (bps < 1) bps = 1;
    aug_gather_kernel<<<N * bps, 256, 0, s>>>(src, tmp, pr, sm, H, W, bps);
    MRISR_LAUNCH_CHECK();
    aug_finish_kernel<<<grid1d(hw * N, 256), 256, 0, s>>>(tmp, dst, dst_u8, pr, sm, noise, hw, hw * N);
    MRISR_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
