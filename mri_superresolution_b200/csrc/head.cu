// Dual-branch head of UNetSuperRes (models/unet_model.py:202-211): blend of the bilinear and the
// PixelShuffle branch with sigmoid(alpha), and the fused GroupNorm+LeakyReLU -> conv1x1(+bias) ->
// sigmoid tail; plus d(alpha) and the per-channel column sum used for conv-bias gradients.
#include <stdlib.h>
#include "kernels.cuh"

namespace mrisr {
namespace {
constexpr int NTHR = 256;

__device__ __forceinline__ float sigmoidf_(float v) { return 1.f / (1.f + __expf(-v)); }

// Per-thread constants of the two branches for one 8-channel vector of the blended tensor: GroupNorm folded into
// scale/shift (z = r * sc + sh).  PixelShuffle(2): shuffled[n, c, 2y+i, 2x+j] = raw[n, 4c+2i+j, y, x]
// (models/unet_model.py:102,111), so output channel c at sub-position ij = 2i+j reads raw channel 4c+ij.
struct BlendCtx {
    float scb[8], shb[8], scp[4][8], shp[4][8];
};
template <typename T>
__device__ __forceinline__ void load_blend_ctx(const BlendArgs<T>& a, int n, int cv, BlendCtx& c) {
    const int gsb = a.r_bil.C / kGroups, gsp = a.r_ps.C / kGroups;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int ch = cv * 8 + k, g = ch / gsb;
        const float mean = a.st_bil[(n * kGroups + g) * 2], rstd = a.st_bil[(n * kGroups + g) * 2 + 1];
        c.scb[k] = rstd * a.g_bil[ch];
        c.shb[k] = a.b_bil[ch] - mean * c.scb[k];
#pragma unroll
        for (int ij = 0; ij < 4; ++ij) {
            const int gp = (4 * ch + ij) / gsp;
            const float mp = a.st_ps[(n * kGroups + gp) * 2], rp = a.st_ps[(n * kGroups + gp) * 2 + 1];
            c.scp[ij][k] = rp * a.g_ps[ch];
            c.shp[ij][k] = a.b_ps[ch] - mp * c.scp[ij][k];
        }
    }
}
// xb / xp of the 2x2 output quad of low-resolution pixel (n, y, x), 8 channels: the quad needs exactly the 32 raw
// pixel-shuffle channels [32 cv, 32 cv + 32) of that pixel (four 8-channel vectors) and four vectors of the bilinear branch
template <typename T>
__device__ __forceinline__ void quad_values(const BlendArgs<T>& a, const BlendCtx& c, int n, int y, int x, int cv,
                                            float (&xb)[4][8], float (&xp)[4][8]) {
    float raw[4][8];
#pragma unroll
    for (int p = 0; p < 4; ++p) load8(a.r_ps.at(n, y, x, (4 * cv + p) * 8), raw[p]);
#pragma unroll
    for (int ij = 0; ij < 4; ++ij) {
        float v[8];
        load8(a.r_bil.at(n, 2 * y + (ij >> 1), 2 * x + (ij & 1), cv * 8), v);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            xb[ij][k] = lrelu(fmaf(v[k], c.scb[k], c.shb[k]));
            xp[ij][k] = lrelu(fmaf(raw[k >> 1][(k & 1) * 4 + ij], c.scp[ij][k], c.shp[ij][k]));
        }
    }
}

// one block = `rows_pb` LOW-resolution rows of one sample; a thread owns one channel vector and walks 2x2 output quads
template <typename T>
__global__ void __launch_bounds__(NTHR, 2) head_blend_kernel(BlendArgs<T> a, int rows_pb) {
    const int n = blockIdx.y;
    const RowMap m = row_map(a.xm, NTHR, threadIdx.x);
    if (!m.active) return;
    const float al = sigmoidf_(a.alpha[0]);
    BlendCtx c;
    load_blend_ctx(a, n, m.cv, c);
    const int h = a.r_ps.H, w = a.r_ps.W;
    const int y_end = min(h, (int)(blockIdx.x + 1) * rows_pb);
    for (int y = blockIdx.x * rows_pb; y < y_end; ++y)
        for (int x = m.x0; x < w; x += m.xstep) {
            float xb[4][8], xp[4][8];
            quad_values(a, c, n, y, x, m.cv, xb, xp);
#pragma unroll
            for (int ij = 0; ij < 4; ++ij) {
                float o[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) o[k] = al * xb[ij][k] + (1.f - al) * xp[ij][k];
                store8(a.xm.at(n, 2 * y + (ij >> 1), 2 * x + (ij & 1), m.cv * 8), o);
            }
        }
}

// d(alpha) from the per-channel sums  sum g * xb  (bilinear branch) and  sum g * xp  (pixel-shuffle
// branch) that gn_bwd's finalize kernels of the two branches leave in `sums` ([2][nch]); one warp
__global__ void dalpha_from_channels_kernel(const float* __restrict__ sums, int nch, const float* __restrict__ alpha,
                                            float* __restrict__ dalpha) {
    double t = 0.0;
    for (int i = threadIdx.x; i < nch; i += 32) t += (double)sums[i] - (double)sums[nch + i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (threadIdx.x == 0) {
        const float s = sigmoidf_(alpha[0]);
        dalpha[0] = (float)(t * (double)(s * (1.f - s)));
    }
}

// one thread per output pixel: all Ch channels (Ch = f/2: 8 or 16 ... up to 64); one block = rows of ONE sample so the
// folded GroupNorm scale/shift and the 1x1 weights sit in shared memory
template <typename T>
__global__ void __launch_bounds__(NTHR) final_head_kernel(Act<T> r, const float* __restrict__ stats,
                                                          const float* __restrict__ gamma,
                                                          const float* __restrict__ beta,
                                                          const float* __restrict__ w1x1,
                                                          const float* __restrict__ b1x1, float* __restrict__ y, int rows_pb) {
    __shared__ float s_sc[256], s_sh[256], s_w[256];
    const int C = r.C, nv = C / 8, gs = C / kGroups;
    const int n = blockIdx.y;
    for (int c = threadIdx.x; c < C; c += NTHR) {
        const int g = c / gs;
        const float mean = stats[(n * kGroups + g) * 2], rstd = stats[(n * kGroups + g) * 2 + 1];
        s_sc[c] = rstd * gamma[c];
        s_sh[c] = beta[c] - mean * s_sc[c];
        s_w[c] = w1x1[c];
    }
    __syncthreads();
    const float b0 = b1x1[0];
    const int y_end = min(r.H, (int)(blockIdx.x + 1) * rows_pb);
    for (int Y = blockIdx.x * rows_pb; Y < y_end; ++Y)
        for (int X = threadIdx.x; X < r.W; X += NTHR) {
            float acc = b0;
            for (int cv = 0; cv < nv; ++cv) {
                float v[8];
                load8(r.at(n, Y, X, cv * 8), v);
#pragma unroll
                for (int k = 0; k < 8; ++k) acc = fmaf(s_w[cv * 8 + k], lrelu(fmaf(v[k], s_sc[cv * 8 + k], s_sh[cv * 8 + k])), acc);
            }
            y[((int64_t)n * r.H + Y) * r.W + X] = 1.f / (1.f + expf(-acc));
        }
}

template <typename T>
__global__ void __launch_bounds__(NTHR) colsum_kernel(Act<T> t, float* __restrict__ scratch, int rows_pb) {
    __shared__ float s_red[NTHR * 8];
    const int C = t.C, nv = C / 8;
    const int tid = threadIdx.x;
    const int n = blockIdx.y;
    const RowMap m = row_map(t, NTHR, tid);
    float s1[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) s1[k] = 0.f;
    if (m.active) {
        const int y_end = min(t.H, (int)(blockIdx.x + 1) * rows_pb);
        for (int y = blockIdx.x * rows_pb; y < y_end; ++y)
            for (int x = m.x0; x < t.W; x += m.xstep) {
                float v[8];
                load8(t.at(n, y, x, m.cv * 8), v);
#pragma unroll
                for (int k = 0; k < 8; ++k) s1[k] += v[k];
            }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) s_red[tid * 8 + k] = s1[k];
    __syncthreads();
    const bool planar = t.planar();
    const int cnt = NTHR / nv;
    for (int c = tid; c < C; c += NTHR) {
        const int cv = c >> 3, k = c & 7;
        float v = 0.f;
        for (int j = 0; j < cnt; ++j) v += s_red[(planar ? cv * cnt + j : cv + j * nv) * 8 + k];
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

}  // namespace

template <typename T>
int launch_head_blend(const BlendArgs<T>& a, cudaStream_t s) {
    ProfScope prof_("head_blend", s);
    // A block first folds both branches' GroupNorm into 80 per-thread constants (160 scalar loads, 40 integer divisions:
    // load_blend_ctx), about as long as two quads of the blend itself (527 instructions per quad): long blocks amortise it.
    constexpr int upt = 8;       // 2x2 quads per thread: swept on B200 in round 2, 2 -> 8: 0.471 -> 0.392 ms per step
    const int rpb = rows_per_block(a.r_ps.W, a.xm.C / 8, NTHR, upt);
    head_blend_kernel<T><<<dim3(ceil_div(a.r_ps.H, rpb), a.xm.N), NTHR, 0, s>>>(a, rpb);
    MRISR_LAUNCH_CHECK();
    return 0;
}

int launch_dalpha_from_channels(const float* sums, int nch, const float* alpha, float* dalpha, cudaStream_t s) {
    ProfScope prof_("blend_dalpha", s);
    dalpha_from_channels_kernel<<<1, 32, 0, s>>>(sums, nch, alpha, dalpha);
    MRISR_LAUNCH_CHECK();
    return 0;
}

template <typename T>
int launch_final_head(Act<T> r, const float* stats, const float* gamma, const float* beta, const float* w1x1,
                      const float* b1x1, float* y, cudaStream_t s) {
    ProfScope prof_("final_head", s);
    if (r.C > 256) { set_error("final_head: C = %d is too wide", r.C); return -1; }
    const int rpb = rows_per_block(r.W, 1, NTHR, 8);
    final_head_kernel<T><<<dim3(ceil_div(r.H, rpb), r.N), NTHR, 0, s>>>(r, stats, gamma, beta, w1x1, b1x1, y, rpb);
    MRISR_LAUNCH_CHECK();
    return 0;
}

constexpr int COLSUM_UPT = 32;
size_t colsum_scratch_elems(int N, int H, int W, int C) {
    return (size_t)N * ceil_div(H, rows_per_block(W, C / 8, NTHR, COLSUM_UPT)) * C;
}

template <typename T>
int launch_colsum(Act<T> t, float* out, float* scratch, cudaStream_t s) {
    ProfScope prof_("colsum", s);
    const int rpb = rows_per_block(t.W, t.C / 8, NTHR, COLSUM_UPT);
    const int slots = ceil_div(t.H, rpb);
    colsum_kernel<T><<<dim3(slots, t.N), NTHR, 0, s>>>(t, scratch, rpb);
    MRISR_LAUNCH_CHECK();
    colsum_finalize_kernel<<<t.C, 32, 0, s>>>(scratch, out, t.N * slots, t.C);
    MRISR_LAUNCH_CHECK();
    return 0;
}

#define INST(T)                                                                                                  \
    template int launch_head_blend<T>(const BlendArgs<T>&, cudaStream_t);                                        \
    template int launch_final_head<T>(Act<T>, const float*, const float*, const float*, const float*,            \
                                      const float*, float*, cudaStream_t);                                       \
    template int launch_colsum<T>(Act<T>, float*, float*, cudaStream_t);
INST(float)
INST(__nv_bfloat16)
#undef INST

}  // namespace mrisr
