// Dual-branch head of UNetSuperRes (models/unet_model.py:202-211): blend of the bilinear and the
// PixelShuffle branch with sigmoid(alpha), and the fused GroupNorm+LeakyReLU -> conv1x1(+bias) ->
// sigmoid tail; plus d(alpha) and the per-channel column sum used for conv-bias gradients.
#include "kernels.cuh"

namespace mrisr {
namespace {
constexpr int NTHR = 256;

__device__ __forceinline__ float sigmoidf_(float v) { return 1.f / (1.f + __expf(-v)); }

// xb / xp (8 channels of the blended-resolution pixel (n, Y, X)) recomputed from the raw conv outputs
template <typename T>
__device__ __forceinline__ void branch_values(const BlendArgs<T>& a, int n, int Y, int X, int cv, float (&xb)[8],
                                              float (&xp)[8]) {
    const int Ch = a.r_bil.C, gsb = Ch / kGroups;          // Ch >= 8
    float v[8];
    load8(a.r_bil.at(n, Y, X, cv * 8), v);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int c = cv * 8 + k, g = c / gsb;
        const float mean = a.st_bil[(n * kGroups + g) * 2], rstd = a.st_bil[(n * kGroups + g) * 2 + 1];
        xb[k] = lrelu((v[k] - mean) * rstd * a.g_bil[c] + a.b_bil[c]);
    }
    // PixelShuffle(2): shuffled[n, c, 2y+i, 2x+j] = raw[n, 4c+2i+j, y, x]   (models/unet_model.py:102,111)
    const int y = Y >> 1, x = X >> 1, ij = ((Y & 1) << 1) | (X & 1);
    const int Craw = a.r_ps.C, gsp = Craw / kGroups;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int c = cv * 8 + k, craw = 4 * c + ij, g = craw / gsp;
        const float mean = a.st_ps[(n * kGroups + g) * 2], rstd = a.st_ps[(n * kGroups + g) * 2 + 1];
        xp[k] = lrelu((to_f(*a.r_ps.at(n, y, x, craw)) - mean) * rstd * a.g_ps[c] + a.b_ps[c]);
    }
}

template <typename T>
__global__ void __launch_bounds__(NTHR) head_blend_kernel(BlendArgs<T> a) {
    const int nv = a.xm.C / 8;
    const float al = sigmoidf_(a.alpha[0]);
    const int64_t total = (int64_t)a.xm.N * a.xm.H * a.xm.W * nv;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int cv = (int)(i % nv);
        const int X = (int)((i / nv) % a.xm.W), Y = (int)((i / ((int64_t)nv * a.xm.W)) % a.xm.H);
        const int n = (int)(i / ((int64_t)nv * a.xm.W * a.xm.H));
        float xb[8], xp[8], o[8];
        branch_values(a, n, Y, X, cv, xb, xp);
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = al * xb[k] + (1.f - al) * xp[k];
        store8(a.xm.at(n, Y, X, cv * 8), o);
    }
}

template <typename T>
__global__ void __launch_bounds__(NTHR) blend_dalpha_kernel(BlendArgs<T> a, Act<T> dxm, float* __restrict__ scratch) {
    __shared__ float s_w[NTHR / 32];
    const int nv = a.xm.C / 8;
    const int64_t total = (int64_t)a.xm.N * a.xm.H * a.xm.W * nv;
    float acc = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int cv = (int)(i % nv);
        const int X = (int)((i / nv) % a.xm.W), Y = (int)((i / ((int64_t)nv * a.xm.W)) % a.xm.H);
        const int n = (int)(i / ((int64_t)nv * a.xm.W * a.xm.H));
        float xb[8], xp[8], g[8];
        branch_values(a, n, Y, X, cv, xb, xp);
        load8(dxm.at(n, Y, X, cv * 8), g);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc = fmaf(g[k], xb[k] - xp[k], acc);
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int w = 0; w < NTHR / 32; ++w) t += s_w[w];
        scratch[blockIdx.x] = t;
    }
}

__global__ void dalpha_finalize_kernel(const float* __restrict__ scratch, int nblocks, const float* __restrict__ alpha,
                                       float* __restrict__ dalpha) {
    double t = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 32) t += (double)scratch[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (threadIdx.x == 0) {
        const float s = sigmoidf_(alpha[0]);
        dalpha[0] = (float)(t * (double)(s * (1.f - s)));
    }
}

// one thread per output pixel: all Ch channels (Ch = f/2: 8 or 16 ... up to 64)
template <typename T>
__global__ void __launch_bounds__(NTHR) final_head_kernel(Act<T> r, const float* __restrict__ stats,
                                                          const float* __restrict__ gamma,
                                                          const float* __restrict__ beta,
                                                          const float* __restrict__ w1x1,
                                                          const float* __restrict__ b1x1, float* __restrict__ y) {
    const int C = r.C, nv = C / 8, gs = C / kGroups;
    const int64_t total = (int64_t)r.N * r.H * r.W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int X = (int)(i % r.W), Y = (int)((i / r.W) % r.H), n = (int)(i / ((int64_t)r.W * r.H));
        float acc = b1x1[0];
        for (int cv = 0; cv < nv; ++cv) {
            float v[8];
            load8(r.at(n, Y, X, cv * 8), v);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int c = cv * 8 + k, g = c / gs;
                const float mean = stats[(n * kGroups + g) * 2], rstd = stats[(n * kGroups + g) * 2 + 1];
                acc = fmaf(w1x1[c], lrelu((v[k] - mean) * rstd * gamma[c] + beta[c]), acc);
            }
        }
        y[i] = 1.f / (1.f + expf(-acc));
    }
}

template <typename T>
__global__ void __launch_bounds__(NTHR) colsum_kernel(Act<T> t, float* __restrict__ scratch, int pix_per_block) {
    __shared__ float s_red[2048];
    const int C = t.C, nv = C / 8, PL = NTHR / nv;
    const int tid = threadIdx.x, cv = tid % nv, pl = tid / nv;
    const int n = blockIdx.y, HW = t.H * t.W;
    float s1[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) s1[k] = 0.f;
    if (pl < PL) {
        const int p_end = min(HW, (int)(blockIdx.x + 1) * pix_per_block);
        for (int p = blockIdx.x * pix_per_block + pl; p < p_end; p += PL) {
            float v[8];
            load8(t.at(n, p / t.W, p % t.W, cv * 8), v);
#pragma unroll
            for (int k = 0; k < 8; ++k) s1[k] += v[k];
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) s_red[pl * C + cv * 8 + k] = s1[k];
    }
    __syncthreads();
    for (int c = tid; c < C; c += NTHR) {
        float v = 0.f;
        for (int j = 0; j < PL; ++j) v += s_red[j * C + c];
        scratch[((int64_t)n * gridDim.x + blockIdx.x) * C + c] = v;
    }
}
__global__ void colsum_finalize_kernel(const float* __restrict__ scratch, float* __restrict__ out, int rows, int C) {
    const int c = blockIdx.x;
    double t = 0.0;
    for (int i = threadIdx.x; i < rows; i += 32) t += (double)scratch[(int64_t)i * C + c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (threadIdx.x == 0) out[c] = (float)t;
}

inline int grid_for(int64_t total) {
    int64_t b = ceil_div64(total, NTHR);
    if (b > 148 * 16) b = 148 * 16;
    if (b < 1) b = 1;
    return (int)b;
}
constexpr int COLSUM_PPT = 32;
}  // namespace

template <typename T>
int launch_head_blend(const BlendArgs<T>& a, cudaStream_t s) {
    ProfScope prof_("head_blend", s);
    head_blend_kernel<T><<<grid_for((int64_t)a.xm.N * a.xm.H * a.xm.W * (a.xm.C / 8)), NTHR, 0, s>>>(a);
    MRISR_LAUNCH_CHECK();
    return 0;
}

size_t blend_dalpha_scratch_elems(int, int, int) { return 148 * 16; }

template <typename T>
int launch_blend_dalpha(const BlendArgs<T>& a, Act<T> dxm, float* dalpha, float* scratch, cudaStream_t s) {
    ProfScope prof_("blend_dalpha", s);
    const int nb = grid_for((int64_t)a.xm.N * a.xm.H * a.xm.W * (a.xm.C / 8));
    blend_dalpha_kernel<T><<<nb, NTHR, 0, s>>>(a, dxm, scratch);
    MRISR_LAUNCH_CHECK();
    dalpha_finalize_kernel<<<1, 32, 0, s>>>(scratch, nb, a.alpha, dalpha);
    MRISR_LAUNCH_CHECK();
    return 0;
}

template <typename T>
int launch_final_head(Act<T> r, const float* stats, const float* gamma, const float* beta, const float* w1x1,
                      const float* b1x1, float* y, cudaStream_t s) {
    ProfScope prof_("final_head", s);
    final_head_kernel<T><<<grid_for((int64_t)r.N * r.H * r.W), NTHR, 0, s>>>(r, stats, gamma, beta, w1x1, b1x1, y);
    MRISR_LAUNCH_CHECK();
    return 0;
}

size_t colsum_scratch_elems(int N, int H, int W, int C) {
    const int PL = NTHR / (C / 8);
    return (size_t)N * ceil_div(H * W, PL * COLSUM_PPT) * C;
}

template <typename T>
int launch_colsum(Act<T> t, float* out, float* scratch, cudaStream_t s) {
    ProfScope prof_("colsum", s);
    const int PL = NTHR / (t.C / 8);
    const int slots = ceil_div(t.H * t.W, PL * COLSUM_PPT);
    colsum_kernel<T><<<dim3(slots, t.N), NTHR, 0, s>>>(t, scratch, PL * COLSUM_PPT);
    MRISR_LAUNCH_CHECK();
    colsum_finalize_kernel<<<t.C, 32, 0, s>>>(scratch, out, t.N * slots, t.C);
    MRISR_LAUNCH_CHECK();
    return 0;
}

#define INST(T)                                                                                                  \
    template int launch_head_blend<T>(const BlendArgs<T>&, cudaStream_t);                                        \
    template int launch_blend_dalpha<T>(const BlendArgs<T>&, Act<T>, float*, float*, cudaStream_t);              \
    template int launch_final_head<T>(Act<T>, const float*, const float*, const float*, const float*,            \
                                      const float*, float*, cudaStream_t);                                       \
    template int launch_colsum<T>(Act<T>, float*, float*, cudaStream_t);
INST(float)
INST(__nv_bfloat16)
#undef INST

}  // namespace mrisr
