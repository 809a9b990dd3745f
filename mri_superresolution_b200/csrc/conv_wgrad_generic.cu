// CUDA-core weight-gradient kernel + weight packing (fp32 parity mode and small-channel layers).
// dW[co][ci][ky][kx] = sum_{n,y,x} in[n, y+ky-1, x+kx-1, ci] * dy[n, y, x, co]
// (autograd of nn.Conv2d in models/unet_model.py; the reference has no hand-written gradient).
#include "kernels.cuh"

namespace mrisr {
namespace {
constexpr int TW = 16, TH = 8, NTHR = 256;
constexpr int CI_T = 8, CO_T = 32;     // one thread per (ci, co) pair, KS*KS accumulators each

template <typename T, int KS>
__global__ void __launch_bounds__(NTHR) wgrad_generic_kernel(WgradArgs<T> a, int splits, int tiles_per_split) {
    constexpr int HALO = KS / 2, PWD = TW + 2 * HALO, PHT = TH + 2 * HALO, TAPS = KS * KS;
    __shared__ float s_in[PHT * PWD][CI_T + 1];
    __shared__ float s_dy[TW * TH][CO_T + 1];
    const int tid = threadIdx.x;
    const int ci_l = tid % CI_T, co_l = tid / CI_T;
    const int H = a.dy.H, W = a.dy.W, Cout = a.dy.C, N = a.dy.N;
    const int C1 = a.src1.C, C2 = a.src2.p ? a.src2.C : 0, Cin = C1 + C2;
    const int ci_tiles = (Cin + CI_T - 1) / CI_T;
    const int ci0 = (blockIdx.x % ci_tiles) * CI_T, co0 = (blockIdx.x / ci_tiles) * CO_T;
    const int tiles_x = (W + TW - 1) / TW, tiles_y = (H + TH - 1) / TH;
    const int total_tiles = N * tiles_x * tiles_y;
    const int split = blockIdx.y;
    float acc[TAPS];
#pragma unroll
    for (int t = 0; t < TAPS; ++t) acc[t] = 0.f;

    const int t_begin = split * tiles_per_split;
    const int t_end = min(total_tiles, t_begin + tiles_per_split);
    for (int tile = t_begin; tile < t_end; ++tile) {
        const int n = tile / (tiles_x * tiles_y);
        const int tr = tile % (tiles_x * tiles_y);
        const int ty0 = (tr / tiles_x) * TH, tx0 = (tr % tiles_x) * TW;
        for (int i = tid; i < PHT * PWD; i += NTHR) {
            const int py = i / PWD, px = i % PWD;
            const int y = ty0 + py - HALO, x = tx0 + px - HALO;
            float v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = 0.f;
            if (y <= H && x <= W) {
                if (Cin >= 8) {
                    if (ci0 < C1) load8(a.src1.at(n, y, x, ci0), v);
                    else load8(a.src2.at(n, y, x, (ci0 - C1)), v);
                } else {
                    for (int k = 0; k < Cin; ++k) v[k] = to_f(*a.src1.at(n, y, x, k));
                }
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) s_in[i][k] = v[k];
        }
        for (int i = tid; i < TW * TH * (CO_T / 8); i += NTHR) {
            const int p = i / (CO_T / 8), cv = i % (CO_T / 8);
            const int y = ty0 + p / TW, x = tx0 + p % TW;
            float v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = 0.f;
            if (y < H && x < W && co0 + cv * 8 < Cout) load8(a.dy.at(n, y, x, co0 + cv * 8), v);
#pragma unroll
            for (int k = 0; k < 8; ++k) s_dy[p][cv * 8 + k] = v[k];
        }
        __syncthreads();
        for (int p = 0; p < TW * TH; ++p) {
            const float d = s_dy[p][co_l];
            const int py = p / TW, px = p % TW;
#pragma unroll
            for (int t = 0; t < TAPS; ++t)
                acc[t] = fmaf(s_in[(py + t / KS) * PWD + px + t % KS][ci_l], d, acc[t]);
        }
        __syncthreads();
    }
    const int ci = ci0 + ci_l, co = co0 + co_l;
    if (ci < Cin && co < Cout) {
        float* o = a.partial + ((int64_t)split * Cout * Cin + (int64_t)co * Cin + ci) * TAPS;
#pragma unroll
        for (int t = 0; t < TAPS; ++t) o[t] = acc[t];
    }
}

__global__ void wgrad_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dw, int64_t n, int splits) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float t = 0.f;
    for (int s = 0; s < splits; ++s) t += partial[(int64_t)s * n + i];
    dw[i] = t;
}

void wgrad_plan(int N, int H, int W, int Cin, int Cout, int& blocks_x, int& splits, int& tiles_per_split) {
    const int ci_tiles = ceil_div(Cin, CI_T), co_tiles = ceil_div(Cout, CO_T);
    blocks_x = ci_tiles * co_tiles;
    const int total_tiles = N * ceil_div(W, TW) * ceil_div(H, TH);
    splits = 1184 / blocks_x;                       // ~8 CTAs per SM overall
    if (splits < 1) splits = 1;
    if (splits > total_tiles) splits = total_tiles;
    tiles_per_split = ceil_div(total_tiles, splits);
    splits = ceil_div(total_tiles, tiles_per_split);
}

template <typename T>
__global__ void pack_weights_kernel(const float* __restrict__ w, T* __restrict__ wf, T* __restrict__ wd, int Cout,
                                    int Cin, int ks) {
    const int taps = ks * ks;
    const int64_t total = (int64_t)Cout * Cin * taps;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int tap = (int)(i % taps);
        const int ci = (int)((i / taps) % Cin);
        const int co = (int)(i / ((int64_t)taps * Cin));
        const float v = w[i];
        wf[((int64_t)tap * Cin + ci) * Cout + co] = from_f<T>(v);
        if (wd) wd[((int64_t)(taps - 1 - tap) * Cout + co) * Cin + ci] = from_f<T>(v);
    }
}
}  // namespace

size_t wgrad_generic_scratch_elems(int N, int H, int W, int Cin, int Cout, int ks) {
    int bx, sp, tps;
    wgrad_plan(N, H, W, Cin, Cout, bx, sp, tps);
    return (size_t)sp * Cout * Cin * ks * ks;
}

template <typename T>
int launch_wgrad_generic(const WgradArgs<T>& a, cudaStream_t s) {
    const char* tag = "wgrad_generic";
    char tbuf[96];
    if (prof_on()) {
        snprintf(tbuf, sizeof(tbuf), "wgrad_generic:k%d:%d>%d@%dx%d", a.ks, a.src1.C + (a.src2.p ? a.src2.C : 0), a.dy.C, a.dy.H, a.dy.W);
        tag = prof_intern(tbuf);
    }
    ProfScope prof_(tag, s);
    const int Cin = a.src1.C + (a.src2.p ? a.src2.C : 0), Cout = a.dy.C;
    int bx, sp, tps;
    wgrad_plan(a.dy.N, a.dy.H, a.dy.W, Cin, Cout, bx, sp, tps);
    dim3 grid(bx, sp);
    if (a.ks == 3) wgrad_generic_kernel<T, 3><<<grid, NTHR, 0, s>>>(a, sp, tps);
    else wgrad_generic_kernel<T, 1><<<grid, NTHR, 0, s>>>(a, sp, tps);
    MRISR_LAUNCH_CHECK();
    const int64_t n = (int64_t)Cout * Cin * a.ks * a.ks;
    wgrad_reduce_kernel<<<(unsigned)ceil_div64(n, 256), 256, 0, s>>>(a.partial, a.dw, n, sp);
    MRISR_LAUNCH_CHECK();
    return 0;
}
template int launch_wgrad_generic<float>(const WgradArgs<float>&, cudaStream_t);
template int launch_wgrad_generic<__nv_bfloat16>(const WgradArgs<__nv_bfloat16>&, cudaStream_t);

template <typename T>
int launch_pack_weights(const float* w, T* wf, T* wd, int Cout, int Cin, int ks, cudaStream_t s) {
    ProfScope prof_("pack_weights", s);
    const int64_t total = (int64_t)Cout * Cin * ks * ks;
    int blocks = (int)ceil_div64(total, 256);
    if (blocks > 1184) blocks = 1184;
    pack_weights_kernel<T><<<blocks, 256, 0, s>>>(w, wf, wd, Cout, Cin, ks);
    MRISR_LAUNCH_CHECK();
    return 0;
}
template int launch_pack_weights<float>(const float*, float*, float*, int, int, int, cudaStream_t);
template int launch_pack_weights<__nv_bfloat16>(const float*, __nv_bfloat16*, __nv_bfloat16*, int, int, int, cudaStream_t);

}  // namespace mrisr
