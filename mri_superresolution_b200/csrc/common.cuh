// Shared device/host helpers for the mrisr sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

namespace mrisr {

constexpr int kGroups = 8;            // GroupNorm(8, C)   models/unet_model.py:30
constexpr float kGnEps = 1e-5f;       // torch.nn.GroupNorm default
constexpr float kSlope = 0.2f;        // LeakyReLU(0.2)    models/unet_model.py:31

void set_error(const char* fmt, ...);
const char* g_err_ptr();

#define MRISR_CUDA_CHECK(expr)                                                              \
    do {                                                                                    \
        cudaError_t _e = (expr);                                                            \
        if (_e != cudaSuccess) {                                                            \
            ::mrisr::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return (int)_e;                                                                 \
        }                                                                                   \
    } while (0)

// every kernel launch goes through this: counts launches (bench.py reports them) and checks for errors
void count_launch();
#define MRISR_LAUNCH_CHECK() do { ::mrisr::count_launch(); MRISR_CUDA_CHECK(cudaGetLastError()); } while (0)

// cudaFuncSetAttribute (shared-memory opt-in, cluster size) is per device: a one-time flag must be kept per device, or a process
// that touches a second GPU launches there without the opt-in.
struct PerDeviceFlag {
    bool f[64] = {};
    bool& cur() { int d = 0; if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= 64) d = 0; return f[d]; }
};

// Optional per-kernel-family timing (mrisr_prof_* in include/mrisr.h): when enabled, each launcher brackets its
// launches with CUDA events on the launching stream; mrisr_prof_report() sums them per tag after a sync.
// Disabled (the default) it costs one predictable branch.
bool prof_on();
const char* prof_intern(const char* tag);      // stable copy of a formatted tag
struct ProfScope {
    int slot;
    cudaStream_t s;
    ProfScope(const char* tag, cudaStream_t stream);
    ~ProfScope();
};

// Padded NHWC activation view: [N][H+2][W+2][C], zero one-pixel border kept as an invariant.
// `p` points at the padded origin (n=0, row -1, col -1, c=0).
//
// Two memory layouts share this view (element (pixel P, channel c) lives at  P * ps + (c >> 3) * cs + (c & 7)):
//   NHWC      ps = C, cs = 8            rows of C channels              (TMA / swizzled K-major tcgen05 kernels, conv_tc.cu)
//   planar-8  ps = 8, cs = 8 * plane    C/8 planes of [pixel][8 ch]     (bulk-copy / no-swizzle tcgen05 kernels, conv_ring.cu)
// where P = n * (H+2)*(W+2) + (y+1)*(W+2) + (x+1) is the padded-linear pixel index and `plane` the pixels per plane
// (all samples plus guard pixels).  ps == 0 (aggregate-initialised views) means NHWC.
template <typename T>
struct Act {
    T* p;
    int N, H, W, C;
    int ps = 0;             // pixel stride in elements (0 -> C)
    int64_t cs = 8;         // stride between 8-channel chunks in elements
    __host__ __device__ __forceinline__ int pw() const { return W + 2; }
    __host__ __device__ __forceinline__ int64_t sample_pixels() const { return (int64_t)(H + 2) * (W + 2); }
    __host__ __device__ __forceinline__ int64_t rows() const { return (int64_t)N * sample_pixels(); }
    __host__ __device__ __forceinline__ bool planar() const { return ps == 8 && C > 8; }
    // padded-linear index of interior pixel (y, x) in [0,H) x [0,W)  (y, x = -1 / H, W address the zero border)
    __host__ __device__ __forceinline__ int64_t pix(int n, int y, int x) const {
        return n * sample_pixels() + (int64_t)(y + 1) * (W + 2) + (x + 1);
    }
    __host__ __device__ __forceinline__ T* at(int n, int y, int x, int c) const {
        return p + pix(n, y, x) * (ps ? ps : C) + (int64_t)(c >> 3) * cs + (c & 7);
    }
};

// planar-8 geometry: pixels of guard on each side of a plane and pixels per plane (rows = N*(H+2)*(W+2))
// (sized so that the ring kernels may fetch whole 256-pixel blocks around every tile without leaving the allocation)
inline int64_t planar_guard(int W) { return (((int64_t)W + 4 + 255) / 256) * 256 + 256; }
inline int64_t planar_plane(int64_t rows, int W) { return ((rows + 255) / 256) * 256 + 2 * planar_guard(W); }

__device__ __forceinline__ float lrelu(float z) { return z > 0.f ? z : kSlope * z; }
__device__ __forceinline__ float lrelu_grad(float z) { return z > 0.f ? 1.f : kSlope; }

__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// 256-bit global store (sm_100+): the 16 bf16 channels of one accumulator chunk, 32-byte aligned
__device__ __forceinline__ void st_global_256(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t e, uint32_t f,
                                              uint32_t g, uint32_t h) {
    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d), "r"(e), "r"(f), "r"(g), "r"(h) : "memory");
}

// `#pragma unroll` for functions that are also compiled for the host (the host pass would warn about the unknown pragma)
#ifdef __CUDA_ARCH__
#define MRISR_UNROLL _Pragma("unroll")
#else
#define MRISR_UNROLL
#endif

// 8-channel vector load/store (16 B for bf16, 32 B for fp32); pointers must be 8-element aligned.
// __host__ __device__: kernel bodies written as plain functions of (block, thread) also run on the CPU in tests/native/.
__host__ __device__ __forceinline__ float bits_to_float(uint32_t u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}
__host__ __device__ __forceinline__ void load8(const float* p, float (&v)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p);
    const float4 b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__host__ __device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 r = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
    for (int i = 0; i < 4; ++i) {
        v[2 * i] = bits_to_float(w[i] << 16);
        v[2 * i + 1] = bits_to_float(w[i] & 0xffff0000u);
    }
}
__host__ __device__ __forceinline__ void store8(float* p, const float (&v)[8]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__host__ __device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
#ifdef __CUDA_ARCH__
    return *reinterpret_cast<uint32_t*>(&t);
#else
    uint32_t u;
    memcpy(&u, &t, 4);
    return u;
#endif
}
__host__ __device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&v)[8]) {
    uint4 r;
    r.x = pack_bf16x2(v[0], v[1]); r.y = pack_bf16x2(v[2], v[3]);
    r.z = pack_bf16x2(v[4], v[5]); r.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(p) = r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Bilinear x2, align_corners=True source coordinates exactly as ATen computes them
// (area_pixel_compute_scale: scale = (in-1)/(out-1) in float; src = scale*dst; i0 = (int)src).
struct Lerp { int i0, i1; float w1; };
__host__ __device__ __forceinline__ float up_scale(int in_size) {
    const int out_size = 2 * in_size;
    return out_size > 1 ? (float)(in_size - 1) / (float)(out_size - 1) : 0.f;
}
__host__ __device__ __forceinline__ Lerp up_coord(int dst, int in_size, float scale) {
    const float s = scale * (float)dst;
    Lerp l;
    l.i0 = (int)s;
    l.i1 = l.i0 + (l.i0 < in_size - 1 ? 1 : 0);
    l.w1 = s - (float)l.i0;
    return l;
}


// Thread -> (8-channel vector, column) mapping of the row-wise elementwise kernels (norm_act.cu, gn_bwd.cu, head.cu).
// Adjacent threads must touch adjacent 16-byte vectors: in a planar-8 tensor those are adjacent PIXELS of one plane, in an
// NHWC tensor adjacent channel vectors of one pixel.  A thread keeps its channel vector for its whole life (so per-channel
// constants and per-channel partial sums live in registers) and walks columns x0, x0 + xstep, ...
struct RowMap { int cv, x0, xstep; bool active; };
template <typename T>
__host__ __device__ __forceinline__ RowMap row_map(const Act<T>& t, int nthr, int tid) {
    const int nv = t.C / 8;
    RowMap m;
    if (t.planar()) {
        const int tpp = nthr / nv;                 // threads per plane
        m.cv = tid / tpp; m.x0 = tid - m.cv * tpp; m.xstep = tpp; m.active = m.cv < nv;
    } else {
        const int pl = nthr / nv;                  // pixels per pass
        m.cv = tid % nv; m.x0 = tid / nv; m.xstep = pl; m.active = m.x0 < pl;
    }
    return m;
}
// rows of `units_w` columns per block so that a block of nthr threads handles about `per_thread` units per thread
inline int rows_per_block(int units_w, int nv, int nthr, int per_thread) {
    if (units_w <= 0 || nv <= 0) return 1;        // e.g. the pooled width of a one-pixel-wide level: no such launch, size queries only
    const int r = (nthr * per_thread) / (units_w * nv);
    return r < 1 ? 1 : r;
}

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

}  // namespace mrisr
