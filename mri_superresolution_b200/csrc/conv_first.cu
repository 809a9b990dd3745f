// First convolution of the network, inc.double_conv.0 (models/unet_model.py:29,137): 3x3, ONE input channel -> f channels.
// K = 9 is not a tensor-core shape and the layer is HBM-bound by its OUTPUT (2 B in, 2*f B out per pixel), so it runs on
// CUDA cores with the planar-8 thread mapping of common.cuh (RowMap): a thread owns 8 output channels, keeps their 72
// weights in registers and walks pixels; GroupNorm partial statistics come out of the same pass.  The weight gradient
//   dW[co][tap] = sum_p X[p + shift(tap)] * dY[p][co]
// uses the same mapping with 72 accumulators per thread and a fixed-order block / grid reduction (deterministic).
// bf16 only (the fp32 parity mode keeps conv_generic.cu); the output / gradient tensor may be planar-8 or NHWC (the layout of
// level 0 at wide images, unet_plan.cuh: level_is_planar) - RowMap gives the thread mapping of either.
#include <stdlib.h>
#include "kernels.cuh"

namespace mrisr {
namespace {
using bf16 = __nv_bfloat16;
constexpr int NTHR = 256;
constexpr int kUnitsBwd = 64;
// output vectors per thread and block of conv_first_kernel.  Swept on B200 in round 2 (64, 256 vectors): 0.199 -> 0.206, 0.217 ms,
// the fixed per-block price (72 scalar weight loads, the statistics reduction) is not what bounds it.
inline int units_fwd() { return 16; }
// rows per block: `units` output vectors per thread, capped so that the staged input rows stay within `tile_bytes`
inline int rows_first(int W, int nv, int units, size_t tile_bytes) {
    int r = rows_per_block(W, nv, NTHR, units);
    const int cap = (int)(tile_bytes / ((size_t)(W + 2) * sizeof(bf16))) - 2;
    if (r > cap) r = cap;
    return r < 1 ? 1 : r;
}
inline int rows_fwd(int W, int nv) { return rows_first(W, nv, units_fwd(), 24 * 1024); }
inline int rows_bwd(int W, int nv) { return rows_first(W, nv, kUnitsBwd, 8 * 1024); }      // + 16 KB ring + 21 KB static < 48 KB

// The input rows a block touches (its rows +- 1, whole padded width: ONE contiguous run of the padded-linear tensor) are staged
// in shared memory first, all loads in flight at once: read per pixel from global memory every column step of every warp waited
// on its own compulsory miss (~1.6 k cycles per step against 135 instructions of work; ncu: 48 % of cycles without an eligible warp).
// x: [N][H+2][W+2][1] (zero border), w: packed [tap][1][Cout] bf16, dst planar-8 [Cout]
__global__ void __launch_bounds__(NTHR) conv_first_kernel(Act<bf16> x, const bf16* __restrict__ w, Act<bf16> dst,
                                                          float* __restrict__ stat_part, int rows_pb) {
    extern __shared__ __align__(16) unsigned char cf_smem[];
    bf16* s_x = reinterpret_cast<bf16*>(cf_smem);               // [(rows + 2)][W + 2]
    __shared__ float s_red[2][NTHR * 8];
    __shared__ float s_ch[2][512];
    const int C = dst.C, nv = C / 8, tid = threadIdx.x, n = blockIdx.y;
    const RowMap m = row_map(dst, NTHR, tid);
    const int pw = x.W + 2;
    const int y_begin = blockIdx.x * rows_pb, y_end = min(dst.H, y_begin + rows_pb);
    {
        const bf16* src = x.p + x.pix(n, y_begin - 1, -1);
        const int count = (y_end - y_begin + 2) * pw;
        for (int i = tid; i < count; i += NTHR) s_x[i] = src[i];
    }
    float s1[8], s2[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) s1[k] = s2[k] = 0.f;
    float wr[9][8];
    if (m.active) {
#pragma unroll
        for (int t = 0; t < 9; ++t)
#pragma unroll
            for (int k = 0; k < 8; ++k) wr[t][k] = to_f(w[t * C + m.cv * 8 + k]);
    }
    __syncthreads();
    if (m.active) {
        const int dstep = m.xstep * (dst.ps ? dst.ps : dst.C);
        for (int y = y_begin; y < y_end; ++y) {
            const bf16* xu = s_x + (y - y_begin) * pw + m.x0 + 1;  // tap (-1, 0) of column x0; the tile starts at (y_begin - 1, -1)
            bf16* dp = dst.at(n, y, m.x0, m.cv * 8);
            for (int xx = m.x0; xx < dst.W; xx += m.xstep, xu += m.xstep, dp += dstep) {
                float v[9];
#pragma unroll
                for (int t = 0; t < 3; ++t) { v[t] = to_f(xu[t - 1]); v[3 + t] = to_f(xu[pw + t - 1]); v[6 + t] = to_f(xu[2 * pw + t - 1]); }
                float o[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    float acc = 0.f;
#pragma unroll
                    for (int t = 0; t < 9; ++t) acc = fmaf(v[t], wr[t][k], acc);
                    o[k] = acc;
                    s1[k] += acc;
                    s2[k] = fmaf(acc, acc, s2[k]);
                }
                store8(dp, o);
            }
        }
    }
    if (!stat_part) return;
#pragma unroll
    for (int k = 0; k < 8; ++k) { s_red[0][tid * 8 + k] = s1[k]; s_red[1][tid * 8 + k] = s2[k]; }
    __syncthreads();
    const int cnt = NTHR / nv;                 // planar mapping: threads [cv*cnt, (cv+1)*cnt) share cv; NHWC: threads cv + j*nv
    const bool planar = dst.planar();
    for (int i = tid; i < 2 * C; i += NTHR) {
        const int which = i / C, c = i % C, cv = c >> 3, k = c & 7;
        float t = 0.f;
        for (int j = 0; j < cnt; ++j) t += s_red[which][(planar ? cv * cnt + j : cv + j * nv) * 8 + k];
        s_ch[which][c] = t;
    }
    __syncthreads();
    if (tid < 2 * kGroups) {
        const int which = tid / kGroups, g = tid % kGroups, gs = C / kGroups;
        float t = 0.f;
        for (int c = g * gs; c < (g + 1) * gs; ++c) t += s_ch[which][c];
        stat_part[(((int64_t)n * gridDim.x + blockIdx.x) * kGroups + g) * 2 + which] = t;
    }
}

// partial[(n * slots + slot)][tap][C] = block sums of X[p + shift(tap)] * dY[p][c]
// Same staging of the input rows; the gradient vectors (16 B per thread and step, streamed once from HBM) run through a
// per-thread cp.async ring kWfDepth steps deep, so a step waits on shared memory instead of a DRAM round trip.
constexpr int kWfDepth = 4;
__device__ __forceinline__ void wf_cp16(void* sdst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(sdst)), "l"(gsrc) : "memory");
}
__global__ void __launch_bounds__(NTHR) wgrad_first_kernel(Act<bf16> x, Act<bf16> dy, float* __restrict__ partial, int rows_pb) {
    extern __shared__ __align__(16) unsigned char cf_smem[];
    uint4* s_g = reinterpret_cast<uint4*>(cf_smem);                         // [kWfDepth][NTHR]
    bf16* s_x = reinterpret_cast<bf16*>(cf_smem + (size_t)kWfDepth * NTHR * 16);   // [(rows + 2)][W + 2]
    __shared__ float s_w[NTHR / 32][72];
    __shared__ float s_wn[NTHR / 32][8][72];
    const int C = dy.C, nv = C / 8, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, n = blockIdx.y;
    const RowMap m = row_map(dy, NTHR, tid);
    const int pw = x.W + 2;
    const int y_begin = blockIdx.x * rows_pb, y_end = min(dy.H, y_begin + rows_pb);
    // this thread's steps: rows y_begin .. y_end - 1, columns x0, x0 + xstep, ... (same count on every row)
    const int per_row = m.active && m.x0 < dy.W ? (dy.W - m.x0 + m.xstep - 1) / m.xstep : 0;
    const int steps = per_row * (y_end - y_begin);
    const int gstep = m.xstep * (dy.ps ? dy.ps : dy.C);
    const int64_t grow = (int64_t)pw * (dy.ps ? dy.ps : dy.C);              // elements between the same column of two rows
    const bf16* gp = m.active ? dy.at(n, y_begin, m.x0, m.cv * 8) : dy.p;   // next vector to fetch
    int fcol = 0;                                                            // its column step inside the row
    auto fetch = [&](int slot) {
        wf_cp16(&s_g[slot * NTHR + tid], gp);
        if (++fcol == per_row) { fcol = 0; gp += grow - (int64_t)(per_row - 1) * gstep; } else gp += gstep;
    };
#pragma unroll
    for (int d = 0; d < kWfDepth - 1; ++d) {
        if (d < steps) fetch(d);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    {
        const bf16* src = x.p + x.pix(n, y_begin - 1, -1);
        const int count = (y_end - y_begin + 2) * pw;
        for (int i = tid; i < count; i += NTHR) s_x[i] = src[i];
    }
    float acc[9][8];
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[t][k] = 0.f;
    __syncthreads();
    {
        int it = 0;
        for (int y = y_begin; y < y_end && per_row; ++y) {
            const bf16* xu = s_x + (y - y_begin) * pw + m.x0 + 1;
            for (int c = 0; c < per_row; ++c, ++it, xu += m.xstep) {
                if (it + kWfDepth - 1 < steps) fetch((it + kWfDepth - 1) % kWfDepth);
                asm volatile("cp.async.commit_group;" ::: "memory");
                asm volatile("cp.async.wait_group %0;" ::"n"(kWfDepth - 1) : "memory");
                const uint4 raw = s_g[(it % kWfDepth) * NTHR + tid];         // written by this thread's own copies only
                const uint32_t r4[4] = {raw.x, raw.y, raw.z, raw.w};
                float g[8], v[9];
#pragma unroll
                for (int i = 0; i < 4; ++i) { g[2 * i] = __uint_as_float(r4[i] << 16); g[2 * i + 1] = __uint_as_float(r4[i] & 0xffff0000u); }
#pragma unroll
                for (int t = 0; t < 3; ++t) { v[t] = to_f(xu[t - 1]); v[3 + t] = to_f(xu[pw + t - 1]); v[6 + t] = to_f(xu[2 * pw + t - 1]); }
#pragma unroll
                for (int t = 0; t < 9; ++t)
#pragma unroll
                    for (int k = 0; k < 8; ++k) acc[t][k] = fmaf(v[t], g[k], acc[t][k]);
            }
        }
    }
    float* out = partial + ((int64_t)n * gridDim.x + blockIdx.x) * 9 * C;
    if (dy.planar()) {
        // a warp's lanes share the channel vector (threads per plane is a multiple of 32): butterfly, then across warps
#pragma unroll
        for (int t = 0; t < 9; ++t)
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float v = acc[t][k];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) s_w[wid][t * 8 + k] = v;
            }
        __syncthreads();
        const int wpc = (NTHR / 32) / nv;                            // warps per channel vector
        for (int i = tid; i < 9 * C; i += NTHR) {
            const int t = i / C, c = i % C, cv = c >> 3, k = c & 7;
            float v = 0.f;
            for (int j = 0; j < wpc; ++j) v += s_w[cv * wpc + j][t * 8 + k];
            out[t * C + c] = v;
        }
    } else {
        // NHWC mapping: lanes l, l + nv, l + 2 nv, ... of every warp share channel vector l (nv divides 32)
#pragma unroll
        for (int t = 0; t < 9; ++t)
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float v = acc[t][k];
                for (int o = 16; o >= nv; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane < nv) s_wn[wid][lane][t * 8 + k] = v;
            }
        __syncthreads();
        for (int i = tid; i < 9 * C; i += NTHR) {
            const int t = i / C, c = i % C, cv = c >> 3, k = c & 7;
            float v = 0.f;
            for (int j = 0; j < NTHR / 32; ++j) v += s_wn[j][cv][t * 8 + k];
            out[t * C + c] = v;
        }
    }
}

// dw[co][0][tap] = sum over blocks of partial[b][tap][co]
__global__ void __launch_bounds__(256) wgrad_first_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dw,
                                                                 int nblocks, int C) {
    __shared__ double s_w[8];
    const int i = blockIdx.x;                                        // tap * C + co
    double t = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += 256) t += (double)partial[(int64_t)b * 9 * C + i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
        double v = 0.0;
        for (int w = 0; w < 8; ++w) v += s_w[w];
        const int tap = i / C, co = i % C;
        dw[co * 9 + tap] = (float)v;
    }
}
}  // namespace

bool conv_first_supported(const Act<__nv_bfloat16>& src, const Act<__nv_bfloat16>& dst, int ks) {
    const int nv = dst.C / 8;
    return ks == 3 && src.C == 1 && dst.C % 8 == 0 && dst.C <= 64 && (NTHR / nv) % 32 == 0 && NTHR % nv == 0 && 32 % nv == 0;
}
int conv_first_stat_slots(int H, int W, int Cout) { return ceil_div(H, rows_fwd(W, Cout / 8)); }
size_t wgrad_first_scratch_elems(int N, int H, int W, int Cout) {
    return (size_t)N * ceil_div(H, rows_bwd(W, Cout / 8)) * 9 * Cout;
}

int launch_conv_first(const ConvArgs<__nv_bfloat16>& a, cudaStream_t s) {
    if (!conv_first_supported(a.src1, a.dst, a.ks) || a.src2.p || a.bias) { set_error("conv_first: unsupported arguments"); return -2; }
    const char* tag = "conv_first";
    char tbuf[96];
    if (prof_on()) {
        snprintf(tbuf, sizeof(tbuf), "conv_first:k3:1>%d@%dx%d", a.dst.C, a.dst.H, a.dst.W);
        tag = prof_intern(tbuf);
    }
    ProfScope prof_(tag, s);
    const int rpb = rows_fwd(a.dst.W, a.dst.C / 8);
    const size_t smem = (size_t)(rpb + 2) * (a.dst.W + 2) * sizeof(__nv_bfloat16);
    static PerDeviceFlag attr_pd; bool& attr_done = attr_pd.cur();
    if (!attr_done) {                                             // very wide images: one staged row triple may pass the 48 KB default
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(conv_first_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        attr_done = true;
    }
    conv_first_kernel<<<dim3(ceil_div(a.dst.H, rpb), a.dst.N), NTHR, smem, s>>>(a.src1, a.w, a.dst, a.stat_part, rpb);
    MRISR_LAUNCH_CHECK();
    return 0;
}

int launch_wgrad_first(const WgradArgs<__nv_bfloat16>& a, cudaStream_t s) {
    if (!conv_first_supported(a.src1, a.dy, a.ks) || a.src2.p) { set_error("wgrad_first: unsupported arguments"); return -2; }
    const char* tag = "wgrad_first";
    char tbuf[96];
    if (prof_on()) {
        snprintf(tbuf, sizeof(tbuf), "wgrad_first:k3:1>%d@%dx%d", a.dy.C, a.dy.H, a.dy.W);
        tag = prof_intern(tbuf);
    }
    ProfScope prof_(tag, s);
    const int rpb = rows_bwd(a.dy.W, a.dy.C / 8);
    const int slots = ceil_div(a.dy.H, rpb);
    const size_t smem = (size_t)kWfDepth * NTHR * 16 + (size_t)(rpb + 2) * (a.dy.W + 2) * sizeof(__nv_bfloat16);
    static PerDeviceFlag attr_pd; bool& attr_done = attr_pd.cur();
    if (!attr_done) {
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(wgrad_first_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        attr_done = true;
    }
    wgrad_first_kernel<<<dim3(slots, a.dy.N), NTHR, smem, s>>>(a.src1, a.dy, a.partial, rpb);
    MRISR_LAUNCH_CHECK();
    wgrad_first_reduce_kernel<<<9 * a.dy.C, 256, 0, s>>>(a.partial, a.dw, slots * a.dy.N, a.dy.C);
    MRISR_LAUNCH_CHECK();
    return 0;
}

}  // namespace mrisr
