// CUDA-core implicit-GEMM convolution (3x3 pad 1 / 1x1) over padded NHWC activations.
// This is the fp32 parity-mode path (MRISR_FP32) and the Cin=1 / tiny-channel path of bf16 mode;
// the bf16 hot layers run on tcgen05 (conv_tc.cu).  Replaces the cuDNN conv calls behind
// nn.Conv2d in models/unet_model.py:29,34,72,101,152,168,172 (fwd, dgrad via flipped weights, wgrad).
#include "kernels.cuh"

namespace mrisr {

namespace {
constexpr int TW = 16, TH = 8, CK = 8, NTHR = 256;

template <typename T, int TC, int KS>
__global__ void __launch_bounds__(NTHR) conv_generic_kernel(ConvArgs<T> a) {
    constexpr int NCT = TC / 4;          // threads along output channels (4 channels each)
    constexpr int NPT = NTHR / NCT;      // threads along pixels
    constexpr int PPT = (TW * TH) / NPT; // pixels per thread
    constexpr int HALO = KS / 2;
    constexpr int PWD = TW + 2 * HALO, PHT = TH + 2 * HALO;
    constexpr int TAPS = KS * KS;
    __shared__ float s_in[PHT * PWD][CK + 1];
    __shared__ __align__(16) float s_w[TAPS][CK][TC];
    __shared__ float s_red[2][NPT][TC];

    const int tid = threadIdx.x;
    const int ct = tid % NCT, pt = tid / NCT;
    const int H = a.dst.H, W = a.dst.W, Cout = a.dst.C;
    const int C1 = a.src1.C, C2 = a.src2.p ? a.src2.C : 0, Cin = C1 + C2;
    const int tiles_x = (W + TW - 1) / TW;
    const int tx0 = (blockIdx.x % tiles_x) * TW, ty0 = (blockIdx.x / tiles_x) * TH;
    const int cbase = blockIdx.y * TC;
    const int n = blockIdx.z;

    float acc[PPT][4];
#pragma unroll
    for (int j = 0; j < PPT; ++j)
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[j][k] = 0.f;

    for (int c0 = 0; c0 < Cin; c0 += CK) {
        // ---- stage the input patch (CK channels) ----
        for (int i = tid; i < PHT * PWD; i += NTHR) {
            const int py = i / PWD, px = i % PWD;
            const int y = ty0 + py - HALO, x = tx0 + px - HALO;   // interior coords, -1..H allowed (zero border)
            float v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = 0.f;
            if (y <= H && x <= W) {
                if (Cin >= 8) {
                    if (c0 < C1) load8(a.src1.at(n, y, x, c0), v);
                    else load8(a.src2.at(n, y, x, (c0 - C1)), v);
                } else {
                    for (int k = 0; k < Cin; ++k) v[k] = to_f(*a.src1.at(n, y, x, k));
                }
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) s_in[i][k] = v[k];
        }
        // ---- stage the weights [tap][ci][co] ----
        for (int i = tid; i < TAPS * CK * TC; i += NTHR) {
            const int co = i % TC, ci = (i / TC) % CK, tap = i / (TC * CK);
            float wv = 0.f;
            if (c0 + ci < Cin) wv = to_f(a.w[((int64_t)tap * Cin + c0 + ci) * Cout + cbase + co]);
            s_w[tap][ci][co] = wv;
        }
        __syncthreads();
#pragma unroll
        for (int tap = 0; tap < TAPS; ++tap) {
            const int dy = tap / KS, dx = tap % KS;
#pragma unroll
            for (int ci = 0; ci < CK; ++ci) {
                const float4 w4 = *reinterpret_cast<const float4*>(&s_w[tap][ci][ct * 4]);
#pragma unroll
                for (int j = 0; j < PPT; ++j) {
                    const int p = pt + j * NPT;
                    const float v = s_in[(p / TW + dy) * PWD + (p % TW) + dx][ci];
                    acc[j][0] = fmaf(v, w4.x, acc[j][0]);
                    acc[j][1] = fmaf(v, w4.y, acc[j][1]);
                    acc[j][2] = fmaf(v, w4.z, acc[j][2]);
                    acc[j][3] = fmaf(v, w4.w, acc[j][3]);
                }
            }
        }
        __syncthreads();
    }

    // ---- epilogue: bias, store, GroupNorm partial statistics ----
    float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
    float b4[4] = {0.f, 0.f, 0.f, 0.f};
    if (a.bias) {
#pragma unroll
        for (int k = 0; k < 4; ++k) b4[k] = a.bias[cbase + ct * 4 + k];
    }
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        const int p = pt + j * NPT;
        const int y = ty0 + p / TW, x = tx0 + p % TW;
        if (y < H && x < W) {
            T* o = a.dst.at(n, y, x, cbase + ct * 4);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float v = acc[j][k] + b4[k];
                o[k] = from_f<T>(v);
                s1[k] += v;
                s2[k] += v * v;
            }
        }
    }
    if (a.stat_part) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            s_red[0][pt][ct * 4 + k] = s1[k];
            s_red[1][pt][ct * 4 + k] = s2[k];
        }
        __syncthreads();
        if (tid < 2 * TC) {                       // fixed-order column sums -> deterministic
            const int which = tid / TC, c = tid % TC;
            float t = 0.f;
            for (int i = 0; i < NPT; ++i) t += s_red[which][i][c];
            s_red[which][0][c] = t;
        }
        __syncthreads();
        if (tid < 2 * kGroups) {
            const int which = tid / kGroups, g = tid % kGroups;
            const int gs = Cout / kGroups;        // channels per group
            float t = 0.f;
            for (int c = 0; c < TC; ++c)
                if ((cbase + c) / gs == g) t += s_red[which][0][c];
            const int slots = gridDim.x * gridDim.y;
            const int slot = blockIdx.x * gridDim.y + blockIdx.y;
            a.stat_part[(((int64_t)n * slots + slot) * kGroups + g) * 2 + which] = t;
        }
    }
}

template <typename T, int TC>
int launch_tc(const ConvArgs<T>& a, cudaStream_t s) {
    const int H = a.dst.H, W = a.dst.W;
    dim3 grid(ceil_div(W, TW) * ceil_div(H, TH), a.dst.C / TC, a.dst.N);
    if (a.ks == 3) conv_generic_kernel<T, TC, 3><<<grid, NTHR, 0, s>>>(a);
    else conv_generic_kernel<T, TC, 1><<<grid, NTHR, 0, s>>>(a);
    MRISR_LAUNCH_CHECK();
    return 0;
}

inline int pick_tc(int Cout) { return Cout % 64 == 0 ? 64 : Cout % 32 == 0 ? 32 : Cout % 16 == 0 ? 16 : 8; }
}  // namespace

int conv_generic_stat_slots(int H, int W, int Cout) {
    return ceil_div(W, TW) * ceil_div(H, TH) * (Cout / pick_tc(Cout));
}

template <typename T>
int launch_conv_generic(const ConvArgs<T>& a, cudaStream_t s) {
    const char* tag = "conv_generic";
    char tbuf[96];
    if (prof_on()) {
        snprintf(tbuf, sizeof(tbuf), "conv_generic:k%d:%d>%d@%dx%d", a.ks, a.src1.C + (a.src2.p ? a.src2.C : 0), a.dst.C, a.dst.H, a.dst.W);
        tag = prof_intern(tbuf);
    }
    ProfScope prof_(tag, s);
    const int Cin = a.src1.C + (a.src2.p ? a.src2.C : 0);
    if (a.dst.C % 8 != 0 || (Cin >= 8 && (a.src1.C % 8 != 0 || Cin % 8 != 0)) || (a.ks != 1 && a.ks != 3)) {
        set_error("conv_generic: unsupported channels Cin=%d (C1=%d) Cout=%d ks=%d", Cin, a.src1.C, a.dst.C, a.ks);
        return -2;
    }
    switch (pick_tc(a.dst.C)) {
        case 64: return launch_tc<T, 64>(a, s);
        case 32: return launch_tc<T, 32>(a, s);
        case 16: return launch_tc<T, 16>(a, s);
        default: return launch_tc<T, 8>(a, s);
    }
}
template int launch_conv_generic<float>(const ConvArgs<float>&, cudaStream_t);
template int launch_conv_generic<__nv_bfloat16>(const ConvArgs<__nv_bfloat16>&, cudaStream_t);

}  // namespace mrisr
