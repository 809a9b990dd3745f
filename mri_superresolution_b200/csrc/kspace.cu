// k-space low-field simulation on the device:
//   utils/preprocessing.py:126-163   robust_normalize: np.percentile(0.5 / 99.5) clip + min-max of a float64 slice
//   utils/preprocessing.py:225-293   simulate_low_field_mri: fft2 -> fftshift -> centre mask -> complex Gaussian noise ->
//                                    ifftshift -> ifft2 -> magnitude -> min-max back to the input's range
//   utils/extraction_utils.py:139-157 LR slice = clip(sim, 0, 1) -> cv2 INTER_AREA to half size -> clip(x * 255) -> uint8
// numpy semantics are the oracle (numpy 2.x: the forward transform of the float32 slice is single precision, the masked
// spectrum, the noise and the inverse transform are complex128).  The transforms are batched cuFFT plans (C2C single forward,
// Z2Z double inverse) working in caller-owned scratch; fftshift / ifftshift are index remaps folded into the one kernel
// between the two transforms (mask + noise), so the shifted spectrum is never materialised.
#include "kernels.cuh"
#include "../../include/mrisr.h"
#include <cub/device/device_segmented_radix_sort.cuh>
#include <cufft.h>
#include <map>
#include <mutex>
#include <tuple>

namespace mrisr {
namespace {
inline int grid1d(long long total, int thr) {
    long long b = (total + thr - 1) / thr;
    if (b > 148 * 32) b = 148 * 32;
    return b < 1 ? 1 : (int)b;
}
inline size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }

#define MRISR_CUFFT_CHECK(expr)                                                                  \
    do {                                                                                         \
        cufftResult _r = (expr);                                                                 \
        if (_r != CUFFT_SUCCESS) { set_error("%s:%d %s -> cufft error %d", __FILE__, __LINE__, #expr, (int)_r); return 1000 + (int)_r; } \
    } while (0)

// one (forward C2C, inverse Z2Z) plan pair per (device, N, H, W); work areas are supplied per call (no allocation by cuFFT)
struct Plans { cufftHandle fwd = 0, inv = 0; size_t ws_fwd = 0, ws_inv = 0; };
int get_plans(int N, int H, int W, Plans& out) {
    static std::mutex mu;
    static std::map<std::tuple<int, int, int, int>, Plans> cache;
    int dev = 0;
    MRISR_CUDA_CHECK(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(mu);
    auto key = std::make_tuple(dev, N, H, W);
    auto it = cache.find(key);
    if (it == cache.end()) {
        Plans p;
        int n[2] = {H, W};
        MRISR_CUFFT_CHECK(cufftCreate(&p.fwd));
        MRISR_CUFFT_CHECK(cufftSetAutoAllocation(p.fwd, 0));
        MRISR_CUFFT_CHECK(cufftMakePlanMany(p.fwd, 2, n, nullptr, 1, H * W, nullptr, 1, H * W, CUFFT_C2C, N, &p.ws_fwd));
        MRISR_CUFFT_CHECK(cufftCreate(&p.inv));
        MRISR_CUFFT_CHECK(cufftSetAutoAllocation(p.inv, 0));
        MRISR_CUFFT_CHECK(cufftMakePlanMany(p.inv, 2, n, nullptr, 1, H * W, nullptr, 1, H * W, CUFFT_Z2Z, N, &p.ws_inv));
        it = cache.emplace(key, p).first;
    }
    out = it->second;
    return 0;
}

struct KsLayout { size_t c64, z128, work, mm, total; };
int ks_layout(int N, int H, int W, KsLayout& L) {
    Plans p;
    const int rc = get_plans(N, H, W, p);
    if (rc) return rc;
    const size_t n = (size_t)N * H * W;
    size_t o = 0;
    L.c64 = o; o += al256(n * sizeof(cufftComplex));
    L.z128 = o; o += al256(n * sizeof(cufftDoubleComplex));
    L.work = o; o += al256(std::max(p.ws_fwd, p.ws_inv));
    L.mm = o; o += al256((size_t)N * 4 * sizeof(double));
    L.total = o;
    return 0;
}

__global__ void __launch_bounds__(256) ks_load_kernel(const float* __restrict__ x, cufftComplex* __restrict__ c, long long total) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x)
        c[i] = make_cuFloatComplex(x[i], 0.f);
}

__device__ __forceinline__ void philox2(long long i, unsigned k0, unsigned k1, unsigned stream_id, double& n0, double& n1) {
    unsigned c0 = (unsigned)i, c1 = (unsigned)((unsigned long long)i >> 32), c2 = stream_id, c3 = 0u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const unsigned hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const unsigned hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    const double u1 = ((double)c0 * 4294967296.0 + (double)c1 + 0.5) * (1.0 / 18446744073709551616.0);
    const double u2 = ((double)c2 * 4294967296.0 + (double)c3 + 0.5) * (1.0 / 18446744073709551616.0);
    const double r = sqrt(-2.0 * log(u1));
    double s, c;
    sincospi(2.0 * u2, &s, &c);
    n0 = r * c; n1 = r * s;
}

// unshifted spectrum index (u, v) <-> shifted index (su, sv) = ((u + H/2) % H, (v + W/2) % W)  (np.fft.fftshift; ifftshift undoes
// it, so the inverse transform consumes the spectrum at its natural index).  Mask and noise are defined on the SHIFTED grid.
__global__ void __launch_bounds__(256) ks_mask_noise_kernel(const cufftComplex* __restrict__ k, cufftDoubleComplex* __restrict__ z,
                                                            const double* __restrict__ noise_re, const double* __restrict__ noise_im,
                                                            int H, int W, int r0, int r1, int c0, int c1, double noise_sd,
                                                            unsigned key0, unsigned key1, long long total) {
    const long long hw = (long long)H * W;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long n = i / hw, p = i % hw;
        const int u = (int)(p / W), v = (int)(p % W);
        const int su = (u + H / 2) % H, sv = (v + W / 2) % W;
        const bool keep = su >= r0 && su < r1 && sv >= c0 && sv < c1;
        const cufftComplex a = k[i];
        double re = keep ? (double)a.x : 0.0, im = keep ? (double)a.y : 0.0;
        const long long si = n * hw + (long long)su * W + sv;
        if (noise_re) { re += noise_re[si]; im += noise_im[si]; }
        else if (noise_sd > 0.0) {
            double e0, e1;
            philox2(si, key0, key1, 0u, e0, e1);
            re += noise_sd * e0; im += noise_sd * e1;
        }
        z[i] = make_cuDoubleComplex(re, im);
    }
}

constexpr int RT = 1024;
// per slice: min / max of |ifft| / (H W) and of the input (orig_min, orig_max), one CTA per slice
__global__ void __launch_bounds__(RT) ks_minmax_kernel(const cufftDoubleComplex* __restrict__ z, const float* __restrict__ x, double* __restrict__ mm,
                                                       long long hw, double inv_hw) {
    __shared__ double s_v[4][RT / 32];
    const cufftDoubleComplex* pz = z + (long long)blockIdx.x * hw;
    const float* px = x + (long long)blockIdx.x * hw;
    double lo = 1e300, hi = -1e300, xlo = 1e300, xhi = -1e300;
    for (long long i = threadIdx.x; i < hw; i += RT) {
        const double m = hypot(pz[i].x * inv_hw, pz[i].y * inv_hw);
        lo = fmin(lo, m); hi = fmax(hi, m);
        const double xv = (double)px[i];
        xlo = fmin(xlo, xv); xhi = fmax(xhi, xv);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        xlo = fmin(xlo, __shfl_xor_sync(0xffffffffu, xlo, o)); xhi = fmax(xhi, __shfl_xor_sync(0xffffffffu, xhi, o));
    }
    if ((threadIdx.x & 31) == 0) { s_v[0][threadIdx.x >> 5] = lo; s_v[1][threadIdx.x >> 5] = hi; s_v[2][threadIdx.x >> 5] = xlo; s_v[3][threadIdx.x >> 5] = xhi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < RT / 32; ++w) { lo = fmin(lo, s_v[0][w]); hi = fmax(hi, s_v[1][w]); xlo = fmin(xlo, s_v[2][w]); xhi = fmax(xhi, s_v[3][w]); }
        double* o = mm + 4 * (long long)blockIdx.x;
        o[0] = lo; o[1] = hi; o[2] = xlo; o[3] = xhi;
    }
}
// simulated = (|z| - min) / (max - min) * (orig_max - orig_min) + orig_min        (preprocessing.py:289-291)
__global__ void __launch_bounds__(256) ks_finish_kernel(const cufftDoubleComplex* __restrict__ z, const double* __restrict__ mm, double* __restrict__ out,
                                                        long long hw, double inv_hw, long long total) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const double* m = mm + 4 * (i / hw);
        const double mag = hypot(z[i].x * inv_hw, z[i].y * inv_hw);
        out[i] = (mag - m[0]) / (m[1] - m[0]) * (m[3] - m[2]) + m[2];
    }
}

// clip(x, 0, 1) -> cv2.resize(INTER_AREA) to exactly half size on float64 (scalar ResizeAreaFast: (s00 + s01 + s10 + s11) * 0.25
// in that order) -> optionally clip(x * 255, 0, 255) truncated to uint8 (extraction_utils.py:143-165)
__global__ void __launch_bounds__(256) area_half_kernel(const double* __restrict__ src, double* __restrict__ dst, uint8_t* __restrict__ dst_u8,
                                                        int H, int W, int clip01, long long total) {
    const int W2 = W / 2, H2 = H / 2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int x = (int)(i % W2);
        const long long r = i / W2;
        const int y = (int)(r % H2);
        const double* p = src + (r / H2) * (long long)H * W + (long long)(2 * y) * W + 2 * x;
        double a = p[0], b = p[1], c = p[W], d = p[W + 1];
        if (clip01) {
            a = fmin(fmax(a, 0.0), 1.0); b = fmin(fmax(b, 0.0), 1.0); c = fmin(fmax(c, 0.0), 1.0); d = fmin(fmax(d, 0.0), 1.0);
        }
        const double v = __dmul_rn(__dadd_rn(__dadd_rn(__dadd_rn(a, b), c), d), 0.25);
        if (dst) dst[i] = v;
        if (dst_u8) {
            const double t = fmin(fmax(__dmul_rn(v, 255.0), 0.0), 255.0);
            dst_u8[i] = (uint8_t)(int)t;
        }
    }
}

// ---- robust_normalize (preprocessing.py:126-163) on float64 slices: order statistics from a segmented sort ---------------
__global__ void rn_offsets_kernel(int* __restrict__ off, int N, int hw) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= N; i += gridDim.x * blockDim.x) off[i] = i * hw;
}
__device__ __forceinline__ double np_percentile_f64(const double* __restrict__ sorted, long long n, double q) {
    const double virt = (double)(n - 1) * (q / 100.0);
    const double fl = floor(virt);
    long long lo = (long long)fl;
    if (lo > n - 1) lo = n - 1;
    const long long hi = lo + 1 < n ? lo + 1 : n - 1;
    const double g = virt - fl, a = sorted[lo], b = sorted[hi], d = b - a;
    double r = __dadd_rn(a, __dmul_rn(d, g));
    if (g >= 0.5) r = __dsub_rn(b, __dmul_rn(d, __dsub_rn(1.0, g)));
    return r;
}
__global__ void __launch_bounds__(256) rn_apply_kernel(const double* __restrict__ x, const double* __restrict__ sorted, float* __restrict__ out,
                                                       double* __restrict__ window, long long hw, double q_lo, double q_hi,
                                                       double t_lo, double t_hi, long long total) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long n = i / hw;
        const double* s = sorted + n * hw;
        const double lower = np_percentile_f64(s, hw, q_lo), upper = np_percentile_f64(s, hw, q_hi);
        if (window && i % hw == 0) { window[2 * n] = lower; window[2 * n + 1] = upper; }
        float r = 0.f;                                           // constant slice, or upper == lower: zeros (:143-153)
        if (s[0] != s[hw - 1] && upper != lower) {
            const double c = fmin(fmax(x[i], lower), upper);
            double v = __ddiv_rn(__dsub_rn(c, lower), __dsub_rn(upper, lower));
            v = __dadd_rn(__dmul_rn(v, __dsub_rn(t_hi, t_lo)), t_lo);
            r = (float)v;
        }
        out[i] = r;
    }
}
struct RnLayout { size_t sorted, off, cub, total; };
RnLayout rn_layout(int N, long long hw) {
    RnLayout L;
    size_t o = 0;
    L.sorted = o; o += al256((size_t)N * hw * sizeof(double));
    L.off = o; o += al256((size_t)(N + 1) * sizeof(int));
    size_t t = 0;
    cub::DeviceSegmentedRadixSort::SortKeys(nullptr, t, (const double*)nullptr, (double*)nullptr, (int)(N * hw), N, (const int*)nullptr, (const int*)nullptr);
    L.cub = o; o += al256(t);
    L.total = o;
    return L;
}
}  // namespace
}  // namespace mrisr

using namespace mrisr;

extern "C" {

size_t mrisr_kspace_scratch_bytes(int N, int H, int W) {
    if (N <= 0 || H <= 0 || W <= 0) return 0;
    KsLayout L;
    if (ks_layout(N, H, W, L)) return 0;
    return L.total;
}

int mrisr_simulate_low_field(const float* data, double* out, const double* noise_real, const double* noise_imag, void* scratch,
                             size_t scratch_bytes, int N, int H, int W, double kspace_crop_factor, double noise_std,
                             uint64_t seed, void* stream) {
    if (!data || !out || !scratch || N <= 0 || H <= 0 || W <= 0 || (noise_real == nullptr) != (noise_imag == nullptr)) { set_error("bad argument"); return MRISR_E_BADARG; }
    KsLayout L;
    int rc = ks_layout(N, H, W, L);
    if (rc) return rc;
    if (scratch_bytes < L.total) { set_error("simulate_low_field: scratch too small: need %zu bytes, got %zu", L.total, scratch_bytes); return MRISR_E_WORKSPACE; }
    Plans p;
    rc = get_plans(N, H, W, p);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    ProfScope prof_("simulate_low_field", s);
    char* b = (char*)scratch;
    auto* c64 = (cufftComplex*)(b + L.c64);
    auto* z = (cufftDoubleComplex*)(b + L.z128);
    double* mm = (double*)(b + L.mm);
    const long long hw = (long long)H * W, total = hw * N;
    // centre block kept by the mask, in shifted coordinates (preprocessing.py:252-266)
    const int crop_r = (int)(H * kspace_crop_factor), crop_c = (int)(W * kspace_crop_factor);
    const int r0 = H / 2 - crop_r / 2, r1 = H / 2 + crop_r / 2, c0 = W / 2 - crop_c / 2, c1 = W / 2 + crop_c / 2;
    const double sd = (noise_std / 255.0) * sqrt((double)H * (double)W) / 10.0;                        // :271
    ks_load_kernel<<<grid1d(total, 256), 256, 0, s>>>(data, c64, total);
    MRISR_LAUNCH_CHECK();
    MRISR_CUFFT_CHECK(cufftSetStream(p.fwd, s));
    MRISR_CUFFT_CHECK(cufftSetWorkArea(p.fwd, b + L.work));
    MRISR_CUFFT_CHECK(cufftExecC2C(p.fwd, c64, c64, CUFFT_FORWARD));
    ks_mask_noise_kernel<<<grid1d(total, 256), 256, 0, s>>>(c64, z, noise_real, noise_imag, H, W, r0, r1, c0, c1, sd,
                                                            (unsigned)seed, (unsigned)(seed >> 32), total);
    MRISR_LAUNCH_CHECK();
    MRISR_CUFFT_CHECK(cufftSetStream(p.inv, s));
    MRISR_CUFFT_CHECK(cufftSetWorkArea(p.inv, b + L.work));
    MRISR_CUFFT_CHECK(cufftExecZ2Z(p.inv, z, z, CUFFT_INVERSE));
    const double inv_hw = 1.0 / ((double)H * (double)W);
    ks_minmax_kernel<<<N, RT, 0, s>>>(z, data, mm, hw, inv_hw);
    MRISR_LAUNCH_CHECK();
    ks_finish_kernel<<<grid1d(total, 256), 256, 0, s>>>(z, mm, out, hw, inv_hw, total);
    MRISR_LAUNCH_CHECK();
    return 0;
}

int mrisr_area_half_f64(const double* src, double* dst, uint8_t* dst_u8, int N, int H, int W, int clip01, void* stream) {
    if (!src || (!dst && !dst_u8) || N <= 0 || H < 2 || W < 2 || (H & 1) || (W & 1)) { set_error("area_half: H and W must be even and >= 2"); return MRISR_E_BADARG; }
    cudaStream_t s = (cudaStream_t)stream;
    ProfScope prof_("area_half", s);
    const long long total = (long long)N * (H / 2) * (W / 2);
    area_half_kernel<<<grid1d(total, 256), 256, 0, s>>>(src, dst, dst_u8, H, W, clip01, total);
    MRISR_LAUNCH_CHECK();
    return 0;
}

size_t mrisr_robust_normalize_scratch_bytes(int N, int H, int W) {
    if (N <= 0 || H <= 0 || W <= 0 || (long long)N * H * W >= (1ll << 31)) return 0;
    return rn_layout(N, (long long)H * W).total;
}

int mrisr_robust_normalize_f64(const double* src, float* dst, double* window, void* scratch, size_t scratch_bytes, int N, int H, int W,
                               double lower_percentile, double upper_percentile, double target_lo, double target_hi, void* stream) {
    if (!src || !dst || !scratch || N <= 0 || H <= 0 || W <= 0) { set_error("bad argument"); return MRISR_E_BADARG; }
    const long long hw = (long long)H * W, total = hw * N;
    if (total >= (1ll << 31)) { set_error("robust_normalize: at most 2^31 - 1 pixels per call"); return MRISR_E_SHAPE; }
    const RnLayout L = rn_layout(N, hw);
    if (scratch_bytes < L.total) { set_error("robust_normalize: scratch too small: need %zu bytes, got %zu", L.total, scratch_bytes); return MRISR_E_WORKSPACE; }
    cudaStream_t s = (cudaStream_t)stream;
    ProfScope prof_("robust_normalize", s);
    char* b = (char*)scratch;
    double* sorted = (double*)(b + L.sorted);
    int* off = (int*)(b + L.off);
    rn_offsets_kernel<<<1, 256, 0, s>>>(off, N, (int)hw);
    MRISR_LAUNCH_CHECK();
    size_t cub_bytes = L.total - L.cub;
    MRISR_CUDA_CHECK(cub::DeviceSegmentedRadixSort::SortKeys(b + L.cub, cub_bytes, src, sorted, (int)total, N, off, off + 1, 0, 64, s));
    rn_apply_kernel<<<grid1d(total, 256), 256, 0, s>>>(src, sorted, dst, window, hw, lower_percentile, upper_percentile, target_lo, target_hi, total);
    MRISR_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
