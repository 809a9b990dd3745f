// Fused L1 + Gaussian-window SSIM loss and gradient (one pass over pred/target, one write of dpred).
// Forward follows utils/losses.py:27-81 (five zero-padded 'same' depthwise Gaussian filters, C1=(0.01R)^2,
// C2=(0.03R)^2, global mean) and :200-226 (l1_w*mean|p-t| + ssim_w*(1-clamp(ssim,0,1))).
// The reference has no hand-written gradient; the analytic form used here (SURVEY.md 8(a) L2'):
//   A1=2*mu1*mu2+C1, A2=2*s12+C2, B1=mu1^2+mu2^2+C1, B2=s11+s22+C2, S=A1*A2/(B1*B2)
//   dE11 = -S/B2;  dE12 = 2*A1/(B1*B2);  dmu1 = 2*mu2*(A2-A1)/(B1*B2) - 2*mu1*S/B1 + 2*mu1*S/B2
//   d mean(S)/d x = ( w*dmu1 + 2x (w*dE11) + y (w*dE12) ) / numel     (w* = same zero-padded filter)
#include "kernels.cuh"
#include <math.h>
#include <stdlib.h>

namespace mrisr {
namespace {
constexpr int T = 32;          // output tile edge
constexpr int NTHR = 320;      // 10 warps: 312 horizontal-pass tasks of the 11-tap training configuration fill them in one round

struct Taps { float w[11]; };

template <int R, bool GRAD>
struct Geo {
    static constexpr int HALO = GRAD ? 2 * R : R;
    static constexpr int IN = T + 2 * HALO;        // staged input edge
    static constexpr int MID = GRAD ? T + 2 * R : T;  // edge of the SSIM-map region
    static constexpr size_t smem_floats = 2 * IN * IN + 5 * IN * MID + (GRAD ? 3 * MID * MID : 0);
};

constexpr int run_len(int extent) { return extent % 8 == 0 ? 8 : (extent % 7 == 0 ? 7 : 8); }

// Register-blocked 1-D filter: a thread produces L consecutive outputs of NM maps from L + WIN - 1 inputs it loads once
// (the plain form reads one shared-memory word per FMA and is LDS-bound; this one reads (L + WIN - 1) / (L * WIN) words).
// Taps are applied in ascending order per output, as in the unblocked form.
template <int L, int NM, int WIN, typename F>
__device__ __forceinline__ void run_filter(const Taps& taps, float (&acc)[L][NM], F&& load) {
#pragma unroll
    for (int o = 0; o < L; ++o)
#pragma unroll
        for (int m = 0; m < NM; ++m) acc[o][m] = 0.f;
#pragma unroll
    for (int j = 0; j < L + WIN - 1; ++j) {
        float p[NM];
        load(j, p);
#pragma unroll
        for (int o = 0; o < L; ++o) {
            const int t = j - o;
            if (t >= 0 && t < WIN) {
#pragma unroll
                for (int m = 0; m < NM; ++m) acc[o][m] = fmaf(taps.w[t], p[m], acc[o][m]);
            }
        }
    }
}

// scratch layout: [nblocks][2] block partials (sum|p-t|, sum S), blocks ordered n-major
template <int R, bool GRAD>
__global__ void __launch_bounds__(NTHR, 2) l1_ssim_kernel(const float* __restrict__ pred, const float* __restrict__ tgt,
                                                          float* __restrict__ dpred, float* __restrict__ partial, int H,
                                                          int W, Taps taps, float C1, float C2, float l1_coef,
                                                          float ssim_coef) {
    using G = Geo<R, GRAD>;
    constexpr int IN = G::IN, MID = G::MID, HALO = G::HALO, WIN = 2 * R + 1;
    extern __shared__ float smem[];
    float* sx = smem;                      // [IN][IN]
    float* sy = sx + IN * IN;              // [IN][IN]
    float* hb = sy + IN * IN;              // [5][IN][MID]  horizontal pass (reused for the backward h-pass)
    float* mp = hb + 5 * IN * MID;         // [3][MID][MID] derivative maps (GRAD only)
    __shared__ float s_red[2][NTHR / 32];

    const int tid = threadIdx.x;
    const int n = blockIdx.z, ty0 = blockIdx.y * T, tx0 = blockIdx.x * T;
    const float* px = pred + (int64_t)n * H * W;
    const float* py = tgt + (int64_t)n * H * W;

    for (int i = tid; i < IN * IN; i += NTHR) {
        const int gy = ty0 - HALO + i / IN, gx = tx0 - HALO + i % IN;
        const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;
        sx[i] = in ? px[(int64_t)gy * W + gx] : 0.f;
        sy[i] = in ? py[(int64_t)gy * W + gx] : 0.f;
    }
    __syncthreads();
    // horizontal pass of x, y, xx, yy, xy: task = (row, run of LH columns)
    {
        constexpr int LH = run_len(MID), RUNS = (MID + LH - 1) / LH;
        for (int task = tid; task < IN * RUNS; task += NTHR) {
            const int r = task / RUNS, c0 = (task % RUNS) * LH;
            float acc[LH][5];
            run_filter<LH, 5, WIN>(taps, acc, [&](int j, float (&p)[5]) {
                const int idx = r * IN + min(c0 + j, IN - 1);
                const float x = sx[idx], y = sy[idx];
                p[0] = x; p[1] = y; p[2] = x * x; p[3] = y * y; p[4] = x * y;
            });
#pragma unroll
            for (int o = 0; o < LH; ++o)
                if (c0 + o < MID) {
#pragma unroll
                    for (int m = 0; m < 5; ++m) hb[m * IN * MID + r * MID + c0 + o] = acc[o][m];
                }
        }
    }
    __syncthreads();
    // vertical pass -> SSIM map (+ derivative maps): task = (run of LV rows, column)
    float ssum = 0.f, lsum = 0.f;
    {
        constexpr int LV = GRAD ? run_len(MID) : 4, RUNS = (MID + LV - 1) / LV;
        for (int task = tid; task < RUNS * MID; task += NTHR) {
            const int r0 = (task / MID) * LV, c = task % MID;
            float acc[LV][5];
            run_filter<LV, 5, WIN>(taps, acc, [&](int j, float (&p)[5]) {
                const int idx = min(r0 + j, IN - 1) * MID + c;
#pragma unroll
                for (int m = 0; m < 5; ++m) p[m] = hb[m * IN * MID + idx];
            });
#pragma unroll
            for (int o = 0; o < LV; ++o) {
                const int r = r0 + o;
                if (r >= MID) continue;
                const float m1 = acc[o][0], m2 = acc[o][1], e11 = acc[o][2], e22 = acc[o][3], e12 = acc[o][4];
                const int off = GRAD ? R : 0;
                const int gy = ty0 - off + r, gx = tx0 - off + c;
                const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;
                const float s11 = e11 - m1 * m1, s22 = e22 - m2 * m2, s12 = e12 - m1 * m2;
                const float A1 = 2.f * m1 * m2 + C1, A2 = 2.f * s12 + C2, B1 = m1 * m1 + m2 * m2 + C1, B2 = s11 + s22 + C2;
                const float inv = 1.f / (B1 * B2);
                const float S = A1 * A2 * inv;
                const bool center = r >= off && r < off + T && c >= off && c < off + T;
                if (in && center) ssum += S;
                if (GRAD) {
                    const float d11 = in ? -S / B2 : 0.f;
                    const float d12 = in ? 2.f * A1 * inv : 0.f;
                    const float d1 = in ? 2.f * m2 * (A2 - A1) * inv - 2.f * m1 * S / B1 + 2.f * m1 * S / B2 : 0.f;
                    const int i = r * MID + c;
                    mp[i] = d1; mp[MID * MID + i] = d11; mp[2 * MID * MID + i] = d12;
                }
            }
        }
    }
    if (GRAD) {
        __syncthreads();
        // horizontal pass of the three derivative maps: rows MID, cols T  (into hb); task = (row, run of 8 columns)
        {
            constexpr int LB = 8, RUNS = T / LB;
            for (int task = tid; task < MID * RUNS; task += NTHR) {
                const int r = task / RUNS, c0 = (task % RUNS) * LB;
                float acc[LB][3];
                run_filter<LB, 3, WIN>(taps, acc, [&](int j, float (&p)[3]) {
                    const int idx = r * MID + c0 + j;
                    p[0] = mp[idx]; p[1] = mp[MID * MID + idx]; p[2] = mp[2 * MID * MID + idx];
                });
#pragma unroll
                for (int o = 0; o < LB; ++o)
#pragma unroll
                    for (int m = 0; m < 3; ++m) hb[m * MID * T + r * T + c0 + o] = acc[o][m];
            }
        }
        __syncthreads();
        // vertical pass + output: task = (run of 4 rows, column)
        {
            constexpr int LB = 4, RUNS = T / LB;
            for (int task = tid; task < RUNS * T; task += NTHR) {
                const int r0 = (task / T) * LB, c = task % T;
                float acc[LB][3];
                run_filter<LB, 3, WIN>(taps, acc, [&](int j, float (&p)[3]) {
                    const int idx = (r0 + j) * T + c;
                    p[0] = hb[idx]; p[1] = hb[MID * T + idx]; p[2] = hb[2 * MID * T + idx];
                });
#pragma unroll
                for (int o = 0; o < LB; ++o) {
                    const int r = r0 + o, gy = ty0 + r, gx = tx0 + c;
                    if (gy < H && gx < W) {
                        const float x = sx[(r + HALO) * IN + c + HALO], y = sy[(r + HALO) * IN + c + HALO];
                        const float d = x - y;
                        lsum += fabsf(d);
                        const float sg = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
                        if (dpred) dpred[(int64_t)n * H * W + (int64_t)gy * W + gx] =
                            l1_coef * sg + ssim_coef * (acc[o][0] + 2.f * x * acc[o][1] + y * acc[o][2]);
                    }
                }
            }
        }
    }
    ssum = warp_sum(ssum);
    lsum = warp_sum(lsum);
    if ((tid & 31) == 0) { s_red[0][tid >> 5] = lsum; s_red[1][tid >> 5] = ssum; }
    __syncthreads();
    if (tid < 2) {
        float t = 0.f;
        for (int w = 0; w < NTHR / 32; ++w) t += s_red[tid][w];
        const int64_t b = ((int64_t)n * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        partial[b * 2 + tid] = t;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Training path (default since round 2; MRISR_SSIM_SPLIT=0 keeps the fused kernel: 0.63 -> 0.57 ms per 64 x 512^2 batch on B200,
// profiles/r02_staged_variants.md): the same loss and gradient as l1_ssim_kernel<R, true> in two kernels with halo R each
// instead of one kernel with halo 2R.  The fused tile recomputes the five forward maps on a
// (T + 2R)^2 region to have the three derivative maps where the backward filters need them: ~430 FMA per output pixel
// against 176 for the separable filters themselves.  Here the forward kernel writes the derivative maps of its OWN pixels
// (d1, d11, d12: 12 B per pixel to scratch) and the gradient kernel filters them with halo R: ~203 FMA per pixel and
// +24 B per pixel of HBM traffic (0.07 ms at batch 64), ~40 KB of shared memory per block instead of 149 KB.
// Every map value is produced by the same run_filter calls in the same tap order as in the fused kernel: the two paths
// agree to fp32 rounding (tests/test_gpu_knobs.py compares a whole train step under MRISR_SSIM_SPLIT=0 with the default).
constexpr int NTHR2 = 256;

template <int R>
__global__ void __launch_bounds__(NTHR2, 4) ssim_fwd_maps_kernel(const float* __restrict__ pred, const float* __restrict__ tgt,
                                                                 float* __restrict__ dmaps, float* __restrict__ partial,
                                                                 int H, int W, Taps taps, float C1, float C2) {
    constexpr int IN = T + 2 * R, WIN = 2 * R + 1;
    extern __shared__ float smem[];
    float* sx = smem;                      // [IN][IN]
    float* sy = sx + IN * IN;              // [IN][IN]
    float* hb = sy + IN * IN;              // [5][IN][T]
    __shared__ float s_red[2][NTHR2 / 32];
    const int tid = threadIdx.x;
    const int n = blockIdx.z, ty0 = blockIdx.y * T, tx0 = blockIdx.x * T;
    const float* px = pred + (int64_t)n * H * W;
    const float* py = tgt + (int64_t)n * H * W;
    for (int i = tid; i < IN * IN; i += NTHR2) {
        const int gy = ty0 - R + i / IN, gx = tx0 - R + i % IN;
        const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;
        sx[i] = in ? px[(int64_t)gy * W + gx] : 0.f;
        sy[i] = in ? py[(int64_t)gy * W + gx] : 0.f;
    }
    __syncthreads();
    {   // horizontal pass of x, y, xx, yy, xy: task = (row, run of 8 columns)
        constexpr int LH = 8, RUNS = T / LH;
        for (int task = tid; task < IN * RUNS; task += NTHR2) {
            const int r = task / RUNS, c0 = (task % RUNS) * LH;
            float acc[LH][5];
            run_filter<LH, 5, WIN>(taps, acc, [&](int j, float (&p)[5]) {
                const int idx = r * IN + c0 + j;
                const float x = sx[idx], y = sy[idx];
                p[0] = x; p[1] = y; p[2] = x * x; p[3] = y * y; p[4] = x * y;
            });
#pragma unroll
            for (int o = 0; o < LH; ++o)
#pragma unroll
                for (int m = 0; m < 5; ++m) hb[m * IN * T + r * T + c0 + o] = acc[o][m];
        }
    }
    __syncthreads();
    float ssum = 0.f, lsum = 0.f;
    {   // vertical pass -> SSIM map and the three derivative maps of the tile's own pixels: task = (run of 4 rows, column)
        constexpr int LV = 4, RUNS = T / LV;
        const int64_t plane = (int64_t)gridDim.z * H * W;
        for (int task = tid; task < RUNS * T; task += NTHR2) {
            const int r0 = (task / T) * LV, c = task % T;
            float acc[LV][5];
            run_filter<LV, 5, WIN>(taps, acc, [&](int j, float (&p)[5]) {
                const int idx = (r0 + j) * T + c;
#pragma unroll
                for (int m = 0; m < 5; ++m) p[m] = hb[m * IN * T + idx];
            });
#pragma unroll
            for (int o = 0; o < LV; ++o) {
                const int r = r0 + o, gy = ty0 + r, gx = tx0 + c;
                if (gy >= H || gx >= W) continue;
                const float m1 = acc[o][0], m2 = acc[o][1], e11 = acc[o][2], e22 = acc[o][3], e12 = acc[o][4];
                const float s11 = e11 - m1 * m1, s22 = e22 - m2 * m2, s12 = e12 - m1 * m2;
                const float A1 = 2.f * m1 * m2 + C1, A2 = 2.f * s12 + C2, B1 = m1 * m1 + m2 * m2 + C1, B2 = s11 + s22 + C2;
                const float inv = 1.f / (B1 * B2);
                const float S = A1 * A2 * inv;
                ssum += S;
                lsum += fabsf(sx[(r + R) * IN + c + R] - sy[(r + R) * IN + c + R]);
                if (dmaps) {
                    const int64_t i = (int64_t)n * H * W + (int64_t)gy * W + gx;
                    dmaps[i] = 2.f * m2 * (A2 - A1) * inv - 2.f * m1 * S / B1 + 2.f * m1 * S / B2;
                    dmaps[plane + i] = -S / B2;
                    dmaps[2 * plane + i] = 2.f * A1 * inv;
                }
            }
        }
    }
    ssum = warp_sum(ssum);
    lsum = warp_sum(lsum);
    if ((tid & 31) == 0) { s_red[0][tid >> 5] = lsum; s_red[1][tid >> 5] = ssum; }
    __syncthreads();
    if (tid < 2) {
        float t = 0.f;
        for (int w = 0; w < NTHR2 / 32; ++w) t += s_red[tid][w];
        const int64_t b = ((int64_t)n * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        partial[b * 2 + tid] = t;
    }
}

template <int R>
__global__ void __launch_bounds__(NTHR2, 4) ssim_grad_kernel(const float* __restrict__ pred, const float* __restrict__ tgt,
                                                             const float* __restrict__ dmaps, float* __restrict__ dpred,
                                                             int H, int W, Taps taps, float l1_coef, float ssim_coef) {
    constexpr int IN = T + 2 * R, WIN = 2 * R + 1;
    extern __shared__ float smem[];
    float* mp = smem;                      // [3][IN][IN] derivative maps, zero outside the image
    float* hb = mp + 3 * IN * IN;          // [3][IN][T]
    const int tid = threadIdx.x;
    const int n = blockIdx.z, ty0 = blockIdx.y * T, tx0 = blockIdx.x * T;
    const int64_t plane = (int64_t)gridDim.z * H * W, base = (int64_t)n * H * W;
    for (int i = tid; i < IN * IN; i += NTHR2) {
        const int gy = ty0 - R + i / IN, gx = tx0 - R + i % IN;
        const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;
        const int64_t g = base + (int64_t)gy * W + gx;
        mp[i] = in ? dmaps[g] : 0.f;
        mp[IN * IN + i] = in ? dmaps[plane + g] : 0.f;
        mp[2 * IN * IN + i] = in ? dmaps[2 * plane + g] : 0.f;
    }
    __syncthreads();
    {   // horizontal pass: task = (row, run of 8 columns)
        constexpr int LB = 8, RUNS = T / LB;
        for (int task = tid; task < IN * RUNS; task += NTHR2) {
            const int r = task / RUNS, c0 = (task % RUNS) * LB;
            float acc[LB][3];
            run_filter<LB, 3, WIN>(taps, acc, [&](int j, float (&p)[3]) {
                const int idx = r * IN + c0 + j;
                p[0] = mp[idx]; p[1] = mp[IN * IN + idx]; p[2] = mp[2 * IN * IN + idx];
            });
#pragma unroll
            for (int o = 0; o < LB; ++o)
#pragma unroll
                for (int m = 0; m < 3; ++m) hb[m * IN * T + r * T + c0 + o] = acc[o][m];
        }
    }
    __syncthreads();
    {   // vertical pass + output: task = (run of 4 rows, column)
        constexpr int LB = 4, RUNS = T / LB;
        for (int task = tid; task < RUNS * T; task += NTHR2) {
            const int r0 = (task / T) * LB, c = task % T;
            float acc[LB][3];
            run_filter<LB, 3, WIN>(taps, acc, [&](int j, float (&p)[3]) {
                const int idx = (r0 + j) * T + c;
                p[0] = hb[idx]; p[1] = hb[IN * T + idx]; p[2] = hb[2 * IN * T + idx];
            });
#pragma unroll
            for (int o = 0; o < LB; ++o) {
                const int gy = ty0 + r0 + o, gx = tx0 + c;
                if (gy < H && gx < W) {
                    const int64_t g = base + (int64_t)gy * W + gx;
                    const float x = pred[g], y = tgt[g];
                    const float d = x - y;
                    const float sg = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
                    dpred[g] = l1_coef * sg + ssim_coef * (acc[o][0] + 2.f * x * acc[o][1] + y * acc[o][2]);
                }
            }
        }
    }
}

template <int R>
int run_split(const float* pred, const float* tgt, float* dpred, float* partial, float* dmaps, int N, int H, int W,
              const Taps& taps, float C1, float C2, float l1_coef, float ssim_coef, cudaStream_t s) {
    constexpr int IN = T + 2 * R;
    const size_t smem_f = (size_t)(2 * IN * IN + 5 * IN * T) * sizeof(float), smem_g = (size_t)(3 * IN * IN + 3 * IN * T) * sizeof(float);
    static PerDeviceFlag configured_pd; bool& configured = configured_pd.cur();
    if (!configured) {
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(ssim_fwd_maps_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_f));
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(ssim_grad_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_g));
        configured = true;
    }
    dim3 grid(ceil_div(W, T), ceil_div(H, T), N);
    ssim_fwd_maps_kernel<R><<<grid, NTHR2, smem_f, s>>>(pred, tgt, dpred ? dmaps : nullptr, partial, H, W, taps, C1, C2);
    MRISR_LAUNCH_CHECK();
    if (dpred) {
        ssim_grad_kernel<R><<<grid, NTHR2, smem_g, s>>>(pred, tgt, dmaps, dpred, H, W, taps, l1_coef, ssim_coef);
        MRISR_LAUNCH_CHECK();
    }
    return 0;
}

int dispatch_split(int win, const float* pred, const float* tgt, float* dpred, float* partial, float* dmaps, int N, int H, int W,
                   const Taps& taps, float C1, float C2, float l1_coef, float ssim_coef, cudaStream_t s) {
    switch (win / 2) {
        case 1: return run_split<1>(pred, tgt, dpred, partial, dmaps, N, H, W, taps, C1, C2, l1_coef, ssim_coef, s);
        case 2: return run_split<2>(pred, tgt, dpred, partial, dmaps, N, H, W, taps, C1, C2, l1_coef, ssim_coef, s);
        case 3: return run_split<3>(pred, tgt, dpred, partial, dmaps, N, H, W, taps, C1, C2, l1_coef, ssim_coef, s);
        case 4: return run_split<4>(pred, tgt, dpred, partial, dmaps, N, H, W, taps, C1, C2, l1_coef, ssim_coef, s);
        default: return run_split<5>(pred, tgt, dpred, partial, dmaps, N, H, W, taps, C1, C2, l1_coef, ssim_coef, s);
    }
}

bool ssim_split_enabled() {
    // on by default: measured on B200 in round 2, 0.627 -> 0.558 ms per 64 x 512^2 batch; MRISR_SSIM_SPLIT=0 = the fused kernel
    static const bool on = [] { const char* e = getenv("MRISR_SSIM_SPLIT"); return !(e && e[0] == '0'); }();
    return on;
}

// out5 = [l1, ssim, total, 0, total]; also a flag in out5-adjacent scratch is not needed: fix-up reads out5[1]
__global__ void loss_finalize_kernel(const float* __restrict__ partial, int nblocks, double inv_numel, float l1_w,
                                     float ssim_w, float* __restrict__ out5) {
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 32) { a += (double)partial[2 * i]; b += (double)partial[2 * i + 1]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
    if (threadIdx.x == 0) {
        const float l1 = (float)(a * inv_numel), ss = (float)(b * inv_numel);
        const float sc = fminf(fmaxf(ss, 0.f), 1.f);
        float total = 0.f;
        if (l1_w > 0.f) total += l1_w * l1;
        if (ssim_w > 0.f) total += ssim_w * (1.f - sc);
        out5[0] = l1; out5[1] = ss; out5[2] = total; out5[3] = 0.f; out5[4] = total;
    }
}

// torch.clamp passes gradient only while 0 <= ssim <= 1 (utils/losses.py:221): when the mean SSIM left that
// range (anti-correlated images) the SSIM part of dpred is dropped.  Rare path: rewrites dpred as L1-only.
__global__ void clamp_fixup_kernel(const float* __restrict__ pred, const float* __restrict__ tgt,
                                   float* __restrict__ dpred, const float* __restrict__ out5, int64_t numel,
                                   float l1_coef) {
    const float ss = out5[1];
    if (ss >= 0.f && ss <= 1.f) return;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel; i += (int64_t)gridDim.x * blockDim.x) {
        const float d = pred[i] - tgt[i];
        dpred[i] = l1_coef * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f));
    }
}

__global__ void ssim_per_sample_kernel(const float* __restrict__ partial, int blocks_per_sample, double inv_hw,
                                       float* __restrict__ out) {
    const int n = blockIdx.x;
    double b = 0.0;
    for (int i = threadIdx.x; i < blocks_per_sample; i += 32) b += (double)partial[2 * ((int64_t)n * blocks_per_sample + i) + 1];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) b += __shfl_xor_sync(0xffffffffu, b, o);
    if (threadIdx.x == 0) out[n] = (float)(b * inv_hw);
}

bool make_taps(int win, float sigma, Taps& t) {
    if (win < 3 || win > 11 || (win & 1) == 0 || !(sigma > 0.f)) return false;
    float g[11], sum = 0.f;                      // utils/losses.py:10-18 in fp32
    for (int i = 0; i < win; ++i) {
        const float c = (float)i - (float)(win / 2);
        g[i] = expf(-(c * c) / (2.f * sigma * sigma));
        sum += g[i];
    }
    for (int i = 0; i < 11; ++i) t.w[i] = i < win ? g[i] / sum : 0.f;
    return true;
}

template <int R, bool GRAD>
int run(const float* pred, const float* tgt, float* dpred, float* partial, int N, int H, int W, const Taps& taps,
        float C1, float C2, float l1_coef, float ssim_coef, cudaStream_t s) {
    const size_t smem = Geo<R, GRAD>::smem_floats * sizeof(float);
    static PerDeviceFlag configured_pd; bool& configured = configured_pd.cur();   // attribute is per-function, idempotent
    if (!configured) {
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(l1_ssim_kernel<R, GRAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = true;
    }
    dim3 grid(ceil_div(W, T), ceil_div(H, T), N);
    l1_ssim_kernel<R, GRAD><<<grid, NTHR, smem, s>>>(pred, tgt, dpred, partial, H, W, taps, C1, C2, l1_coef, ssim_coef);
    MRISR_LAUNCH_CHECK();
    return 0;
}

template <bool GRAD>
int dispatch(int win, const float* pred, const float* tgt, float* dpred, float* partial, int N, int H, int W,
             const Taps& taps, float C1, float C2, float l1_coef, float ssim_coef, cudaStream_t s) {
    switch (win / 2) {
        case 1: return run<1, GRAD>(pred, tgt, dpred, partial, N, H, W, taps, C1, C2, l1_coef, ssim_coef, s);
        case 2: return run<2, GRAD>(pred, tgt, dpred, partial, N, H, W, taps, C1, C2, l1_coef, ssim_coef, s);
        case 3: return run<3, GRAD>(pred, tgt, dpred, partial, N, H, W, taps, C1, C2, l1_coef, ssim_coef, s);
        case 4: return run<4, GRAD>(pred, tgt, dpred, partial, N, H, W, taps, C1, C2, l1_coef, ssim_coef, s);
        default: return run<5, GRAD>(pred, tgt, dpred, partial, N, H, W, taps, C1, C2, l1_coef, ssim_coef, s);
    }
}
}  // namespace

static size_t partial_bytes(int N, int H, int W) {
    return (((size_t)N * ceil_div(H, T) * ceil_div(W, T) * 2 * sizeof(float)) + 255) & ~(size_t)255;
}
size_t l1_ssim_scratch_bytes(int N, int H, int W) {
    // block partials, and (two-kernel training path) the three derivative maps
    return partial_bytes(N, H, W) + (ssim_split_enabled() ? (size_t)3 * N * H * W * sizeof(float) : 0);
}

int launch_l1_ssim(const float* pred, const float* target, float* dpred, float* out5, void* scratch, int N, int H,
                   int W, float l1_w, float ssim_w, int win, float sigma, float val_range, cudaStream_t s) {
    ProfScope prof_("l1_ssim", s);
    Taps taps;
    if (!make_taps(win, sigma, taps)) { set_error("l1_ssim: window_size must be odd in [3,11], sigma > 0"); return -1; }
    const float C1 = (0.01f * val_range) * (0.01f * val_range), C2 = (0.03f * val_range) * (0.03f * val_range);
    const double numel = (double)N * H * W;
    const float l1_coef = l1_w > 0.f ? (float)(l1_w / numel) : 0.f;
    const float ssim_coef = ssim_w > 0.f ? (float)(-ssim_w / numel) : 0.f;
    float* partial = (float*)scratch;
    int rc;
    const int nblocks = N * ceil_div(H, T) * ceil_div(W, T);
    if (ssim_split_enabled()) {
        float* dmaps = (float*)((char*)scratch + partial_bytes(N, H, W));
        rc = dispatch_split(win, pred, target, dpred, partial, dmaps, N, H, W, taps, C1, C2, l1_coef, ssim_coef, s);
    } else {
        rc = dispatch<true>(win, pred, target, dpred, partial, N, H, W, taps, C1, C2, l1_coef, ssim_coef, s);
    }
    if (rc) return rc;
    loss_finalize_kernel<<<1, 32, 0, s>>>(partial, nblocks, 1.0 / numel, l1_w, ssim_w, out5);
    MRISR_LAUNCH_CHECK();
    if (dpred && ssim_w > 0.f) {
        clamp_fixup_kernel<<<148 * 4, 256, 0, s>>>(pred, target, dpred, out5, (int64_t)N * H * W, l1_coef);
        MRISR_LAUNCH_CHECK();
    }
    return 0;
}

int launch_ssim_forward(const float* a, const float* b, float* per_sample, void* scratch, int N, int H, int W,
                        int win, float sigma, float val_range, cudaStream_t s) {
    ProfScope prof_("ssim_forward", s);
    Taps taps;
    if (!make_taps(win, sigma, taps)) { set_error("ssim: window_size must be odd in [3,11], sigma > 0"); return -1; }
    const float C1 = (0.01f * val_range) * (0.01f * val_range), C2 = (0.03f * val_range) * (0.03f * val_range);
    float* partial = (float*)scratch;
    int rc = dispatch<false>(win, a, b, nullptr, partial, N, H, W, taps, C1, C2, 0.f, 0.f, s);
    if (rc) return rc;
    ssim_per_sample_kernel<<<N, 32, 0, s>>>(partial, ceil_div(H, T) * ceil_div(W, T), 1.0 / ((double)H * W), per_sample);
    MRISR_LAUNCH_CHECK();
    return 0;
}

}  // namespace mrisr

// ---------------------------------------------------------------------------------------------------------------------
// Allocation-free SSIM metric (no counterpart in the reference; added for callers that
// cannot allocate: code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same kernels as
// launch_ssim_forward; the block partials go to a file-scope device array instead of caller scratch.  Every GPU that loads
// the library holds the array (MRISR_SSIM_STATIC_SCRATCH_BYTES = 1 MiB) until the process exits.
namespace mrisr {
namespace {
constexpr size_t kStaticScratchBytes = 1u << 20;      // 131 072 tiles of 32 x 32 pixels, e.g. 128 slices of 1024 x 1024
static __device__ __align__(256) unsigned char g_ssim_static_scratch[kStaticScratchBytes];
constexpr int kMaxDevices = 64;

// In host code `g_ssim_static_scratch` is a host-side stand-in, not a device address: the device address is looked up once per
// device, with that device current, and cached.
int static_scratch_ptr(float** out) {
    static void* ptrs[kMaxDevices] = {};
    int dev = -1;
    MRISR_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) { set_error("ssim_forward_static: device index %d out of range", dev); return -1; }
    if (!ptrs[dev]) {
        void* p = nullptr;
        MRISR_CUDA_CHECK(cudaGetSymbolAddress(&p, g_ssim_static_scratch));
        if (!p) { set_error("ssim_forward_static: cudaGetSymbolAddress returned a null address"); return -1; }
        ptrs[dev] = p;
    }
    *out = (float*)ptrs[dev];
    return 0;
}
}  // namespace

size_t ssim_static_scratch_capacity() { return kStaticScratchBytes; }

int launch_ssim_forward_static(const float* a, const float* b, float* per_sample, int N, int H, int W, int win, float sigma,
                               float val_range, cudaStream_t s) {
    const size_t need = partial_bytes(N, H, W);
    if (need > kStaticScratchBytes) {
        set_error("ssim_forward_static: N=%d H=%d W=%d needs %zu bytes of block partials, the static scratch holds %zu "
                  "(use mrisr_ssim_forward with caller-owned scratch)", N, H, W, need, kStaticScratchBytes);
        return -2;
    }
    float* partial = nullptr;
    const int rc = static_scratch_ptr(&partial);
    if (rc) return rc;
    // nothing has to start at zero: every block of the launch overwrites its own (sum|a-b|, sum S) pair before the per-sample
    // reduction reads exactly those pairs; whatever an earlier call left beyond them is never read
    return launch_ssim_forward(a, b, per_sample, partial, N, H, W, win, sigma, val_range, s);
}
}  // namespace mrisr

extern "C" int mrisr_ssim_forward_static(const float* img1, const float* img2, float* per_sample, int N, int H, int W,
                                         int window_size, float sigma, float val_range, void* stream) {
    using namespace mrisr;
    if (!img1 || !img2 || !per_sample || N <= 0 || H <= 0 || W <= 0) { set_error("bad argument"); return -1; }
    return launch_ssim_forward_static(img1, img2, per_sample, N, H, W, window_size, sigma, val_range, (cudaStream_t)stream);
}

extern "C" size_t mrisr_ssim_static_scratch_bytes(void) { return mrisr::ssim_static_scratch_capacity(); }
