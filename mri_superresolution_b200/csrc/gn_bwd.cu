// Backward of conv -> GroupNorm(8) -> LeakyReLU(0.2) blocks: given a description of the gradient
// wrt the activation ("source": plain tensor, + max-pool routing, rank-1 final-conv form, blend /
// pixel-shuffle adjoints) produce the gradient wrt the raw conv output, d(gamma), d(beta).
// The reference has no hand-written backward (autograd of models/unet_model.py:29-37,52,93,110-113,
// 206-211); formulas: z = gamma*xh + beta, xh = (r-mean)*rstd, dz = dA*lrelu'(z),
//   dr = rstd * (gamma*dz - mean_g(gamma*dz) - xh * mean_g(gamma*dz*xh)).
// Kernel form (fewer per-thread constants, more loads in flight): with sc = gamma*rstd, sh = beta - mean*sc the reduce
// pass accumulates per channel  s1 = sum dz,  s2r = sum dz*r  (raw r);  the finalize forms  sum dz*xh = rstd*(s2r - mean*s1),
// m1, m2 and the per-(n, group) constants  q = -rstd*m1 + mean*rstd^2*m2,  p = -rstd^2*m2,  and the apply pass is
//   dr = sc*dz + p*r + q.
#include <stdlib.h>
#include "kernels.cuh"

namespace mrisr {
namespace {
constexpr int NTHR = 256;
constexpr int UNITS_PER_THREAD = 32;
constexpr int FTHR = 256;         // threads of the finalize kernel

// Optional chunking of the batch (MRISR_GNBWD_CHUNK_MB > 0): reduce + apply run per chunk of samples whose (r, g, dr) fit
// the given size, hoping for L2 hits in the apply pass.  Measured on B200 (batch 64, profiles/r01_gn_bwd_chunking.md): every
// chunk size was SLOWER than one pass over the whole batch (launch gaps and tails of the short kernels outweigh the saved
// HBM reads), so the default is off; the knob stays for experiments.
inline double gn_bwd_chunk_bytes() {
    static double v = -1.0;
    if (v < 0.0) {
        const char* e = getenv("MRISR_GNBWD_CHUNK_MB");
        v = e ? atof(e) * 1048576.0 : 0.0;
        if (v <= 0.0) v = 1e15;
    }
    return v;
}

template <int KIND> struct KindTraits { static constexpr bool pool = (KIND == SRC_DIRECT_POOL || KIND == SRC_POOL_ONLY); };

// Per-thread context for 8 channels of one sample: GroupNorm folded into z = r * sc + sh
struct ChanCtx {
    float sc[8], sh[8];
};

template <typename T, int KIND>
__device__ __forceinline__ void load_ctx(const GnBwdArgs<T>& a, int n, int cv, ChanCtx& c) {
    const int gs = a.r.C / kGroups;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int ch = cv * 8 + k, g = ch / gs;
        const int pc = KIND == SRC_BLEND_PS ? ch / 4 : ch;
        const float mean = a.stats[(n * kGroups + g) * 2], rstd = a.stats[(n * kGroups + g) * 2 + 1];
        c.sc[k] = a.gamma[pc] * rstd;
        c.sh[k] = a.beta[pc] - mean * c.sc[k];
    }
}

// Loads the raw conv outputs rv and computes dz for one unit (1 pixel, or a 2x2 window for the pool kinds).
// dl_out: (SRC_OUTER) dl of the pixel; act_out: lrelu(z) (SRC_OUTER weight gradient)
template <typename T, int KIND, int P>
__device__ __forceinline__ void unit_grads(const GnBwdArgs<T>& a, const ChanCtx& c, int n, int uy, int ux, int cv,
                                           float (&rv)[P][8], float (&dz)[P][8], float& dl_out, float (&act_out)[8]) {
    float z[P][8];
#pragma unroll
    for (int q = 0; q < P; ++q) {
        const int y = P == 4 ? 2 * uy + (q >> 1) : uy, x = P == 4 ? 2 * ux + (q & 1) : ux;
        load8(a.r.at(n, y, x, cv * 8), rv[q]);
#pragma unroll
        for (int k = 0; k < 8; ++k) z[q][k] = fmaf(rv[q][k], c.sc[k], c.sh[k]);
        float dA[8];
        if (KIND == SRC_DIRECT || KIND == SRC_DIRECT_POOL) {
            load8(a.g.at(n, y, x, a.coff + cv * 8), dA);
        } else if (KIND == SRC_BLEND_BIL) {
            const float al = 1.f / (1.f + __expf(-a.alpha[0]));
            load8(a.g.at(n, y, x, cv * 8), dA);
#pragma unroll
            for (int k = 0; k < 8; ++k) dA[k] *= al;
        } else if (KIND == SRC_BLEND_PS) {
            const float om = 1.f - 1.f / (1.f + __expf(-a.alpha[0]));
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int ch = cv * 8 + k, cs = ch >> 2, i = (ch >> 1) & 1, j = ch & 1;
                dA[k] = om * to_f(*a.g.at(n, 2 * y + i, 2 * x + j, cs));
            }
        } else if (KIND == SRC_OUTER) {
            const int64_t pi = ((int64_t)n * a.r.H + y) * a.r.W + x;
            const float yy = a.y[pi];
            const float dl = a.dy[pi] * yy * (1.f - yy);
            dl_out = dl;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                dA[k] = dl * a.wvec[cv * 8 + k];
                act_out[k] = lrelu(z[q][k]);
            }
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) dA[k] = 0.f;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) dz[q][k] = dA[k];
    }
    if (KindTraits<KIND>::pool) {
        // route the pooled gradient to the first maximum in scan order (ATen max_pool2d uses '>'),
        // comparing the activations as stored by the forward pass (rounded to T)
        float gp[8];
        load8(a.gpool.at(n, uy, ux, cv * 8), gp);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            int best = 0;
            float bv = to_f(from_f<T>(lrelu(z[0][k])));
#pragma unroll
            for (int q = 1; q < P; ++q) {
                const float vq = to_f(from_f<T>(lrelu(z[q][k])));
                if (vq > bv) { bv = vq; best = q; }
            }
#pragma unroll
            for (int q = 0; q < P; ++q) dz[q][k] += (q == best) ? gp[k] : 0.f;
        }
    }
#pragma unroll
    for (int q = 0; q < P; ++q)
#pragma unroll
        for (int k = 0; k < 8; ++k) dz[q][k] *= lrelu_grad(z[q][k]);
}

// part layout: [N][slots][3][C+8]  (row 0: sum dz, row 1: sum dz*r (raw), row 2: SRC_OUTER sum dl*act; column C: sum dl)
// One block = `rows_pb` unit rows of one sample; thread mapping per common.cuh RowMap (coalesced for both layouts).
template <typename T, int KIND>
__global__ void __launch_bounds__(NTHR) gn_bwd_reduce_kernel(GnBwdArgs<T> a, int rows_pb, int n0) {
    constexpr int P = KindTraits<KIND>::pool ? 4 : 1;
    __shared__ float s_red[3][NTHR * 8];
    __shared__ float s_dl[NTHR];
    const int C = a.r.C, nv = C / 8;
    const int tid = threadIdx.x;
    const int n = n0 + blockIdx.y;
    const int UW = P == 4 ? a.r.W / 2 : a.r.W, UH = P == 4 ? a.r.H / 2 : a.r.H;
    const RowMap m = row_map(a.r, NTHR, tid);
    float s1[8], s2[8], s3[8], sdl = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s1[k] = s2[k] = s3[k] = 0.f;
    if (m.active) {
        ChanCtx c;
        load_ctx<T, KIND>(a, n, m.cv, c);
        const int uy_end = min(UH, (int)(blockIdx.x + 1) * rows_pb);
        constexpr int UB = P == 4 ? 1 : 2;             // units in flight per thread (all loads issued before any use)
        for (int uy = blockIdx.x * rows_pb; uy < uy_end; ++uy) {
            for (int ux0 = m.x0; ux0 < UW; ux0 += UB * m.xstep) {
                float rv[UB][P][8], dz[UB][P][8], act[UB][8], dl[UB];
#pragma unroll
                for (int u = 0; u < UB; ++u) {
                    dl[u] = 0.f;
                    const int ux = min(ux0 + u * m.xstep, UW - 1);      // clamped duplicate is discarded below
                    unit_grads<T, KIND, P>(a, c, n, uy, ux, m.cv, rv[u], dz[u], dl[u], act[u]);
                }
#pragma unroll
                for (int u = 0; u < UB; ++u) {
                    if (ux0 + u * m.xstep >= UW) continue;
#pragma unroll
                    for (int q = 0; q < P; ++q)
#pragma unroll
                        for (int k = 0; k < 8; ++k) { s1[k] += dz[u][q][k]; s2[k] = fmaf(dz[u][q][k], rv[u][q][k], s2[k]); }
                    if (KIND == SRC_OUTER) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) s3[k] += dl[u] * act[u][k];
                        if (m.cv == 0) sdl += dl[u];
                    }
                }
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        s_red[0][tid * 8 + k] = s1[k];
        s_red[1][tid * 8 + k] = s2[k];
        s_red[2][tid * 8 + k] = s3[k];
    }
    s_dl[tid] = sdl;
    __syncthreads();
    const int64_t row = (int64_t)(C + 8);
    float* out = a.part + ((int64_t)n * gridDim.x + blockIdx.x) * 3 * row;
    const bool planar = a.r.planar();
    const int cnt = NTHR / nv;                       // threads sharing one channel vector
    for (int i = tid; i < 3 * C; i += NTHR) {
        const int which = i / C, ch = i % C, cv = ch >> 3, k = ch & 7;
        float t = 0.f;
        for (int j = 0; j < cnt; ++j) t += s_red[which][(planar ? cv * cnt + j : cv + j * nv) * 8 + k];
        out[which * row + ch] = t;
    }
    if (KIND == SRC_OUTER && tid == 0) {
        float t = 0.f;
        for (int j = 0; j < NTHR; ++j) t += s_dl[j];
        out[C] = t;
    }
}

// blocks [0, N*8): the apply constants (q, p) per (n, group).  blocks [N*8, N*8 + C): per-channel parameter gradients.
template <typename T, int KIND>
__global__ void __launch_bounds__(FTHR) gn_bwd_finalize_kernel(GnBwdArgs<T> a, int slots, double inv_count, int n0, int nn, int do_params) {
    __shared__ double s_w[4][FTHR / 32];
    const int C = a.r.C, gs = C / kGroups, N = a.r.N;
    const int64_t row = (int64_t)(C + 8);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, tid = threadIdx.x;
    if (!do_params) {
        const int n = n0 + blockIdx.x / kGroups, g = blockIdx.x % kGroups;
        const int ng = n * kGroups + g;
        if (n >= n0 + nn) return;
        const double mean = (double)a.stats[ng * 2], rstd = (double)a.stats[ng * 2 + 1];
        double m1 = 0.0, m2 = 0.0;
        for (int i = tid; i < slots * gs; i += FTHR) {
            const int s = i / gs, ch = g * gs + i % gs;
            const float* q = a.part + ((int64_t)n * slots + s) * 3 * row;
            const double gm = (double)a.gamma[KIND == SRC_BLEND_PS ? ch / 4 : ch];
            m1 += gm * (double)q[ch];
            m2 += gm * ((double)q[row + ch] - mean * (double)q[ch]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m1 += __shfl_xor_sync(0xffffffffu, m1, o);
            m2 += __shfl_xor_sync(0xffffffffu, m2, o);
        }
        if (lane == 0) { s_w[0][wid] = m1; s_w[1][wid] = m2; }
        __syncthreads();
        if (tid == 0) {
            m1 = m2 = 0.0;
            for (int w = 0; w < FTHR / 32; ++w) { m1 += s_w[0][w]; m2 += s_w[1][w]; }
            m1 *= inv_count;
            m2 *= rstd * inv_count;                       // mean_g(gamma * dz * xh)
            a.mstat[ng * 2] = (float)(-rstd * m1 + mean * rstd * rstd * m2);      // q
            a.mstat[ng * 2 + 1] = (float)(-rstd * rstd * m2);                      // p
        }
    } else {
        // one block per parameter channel pc; raw channels mapping to it: [pc*rep, (pc+1)*rep)
        const int rep = KIND == SRC_BLEND_PS ? 4 : 1;
        const int pc = blockIdx.x;
        double d1 = 0.0, d2 = 0.0, d3 = 0.0, d4 = 0.0;
        for (int i = tid; i < N * slots * rep; i += FTHR) {
            const int ch = pc * rep + i % rep;
            const int ns = i / rep, n = ns / slots, g = ch / gs;
            const float* q = a.part + (int64_t)ns * 3 * row;
            const double mean = (double)a.stats[(n * kGroups + g) * 2], rstd = (double)a.stats[(n * kGroups + g) * 2 + 1];
            d1 += (double)q[ch];
            d2 += rstd * ((double)q[row + ch] - mean * (double)q[ch]);
            if (KIND == SRC_OUTER) {
                d3 += (double)q[2 * row + ch];
                if (pc == 0) d4 += (double)q[C];
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            d1 += __shfl_xor_sync(0xffffffffu, d1, o);
            d2 += __shfl_xor_sync(0xffffffffu, d2, o);
            d3 += __shfl_xor_sync(0xffffffffu, d3, o);
            d4 += __shfl_xor_sync(0xffffffffu, d4, o);
        }
        if (lane == 0) { s_w[0][wid] = d1; s_w[1][wid] = d2; s_w[2][wid] = d3; s_w[3][wid] = d4; }
        __syncthreads();
        if (tid == 0) {
            d1 = d2 = d3 = d4 = 0.0;
            for (int w = 0; w < FTHR / 32; ++w) { d1 += s_w[0][w]; d2 += s_w[1][w]; d3 += s_w[2][w]; d4 += s_w[3][w]; }
            a.dbeta[pc] = (float)d1;
            a.dgamma[pc] = (float)d2;
            if (KIND == SRC_OUTER) {
                a.dwvec[pc] = (float)d3;
                if (pc == 0) a.dbias[0] = (float)d4;
            }
        }
    }
}

template <typename T, int KIND>
__global__ void __launch_bounds__(NTHR) gn_bwd_apply_kernel(GnBwdArgs<T> a, int rows_pb, int n0) {
    constexpr int P = KindTraits<KIND>::pool ? 4 : 1;
    const int C = a.r.C, gs = C / kGroups;
    const int UW = P == 4 ? a.r.W / 2 : a.r.W, UH = P == 4 ? a.r.H / 2 : a.r.H;
    const int n = n0 + blockIdx.y;
    const RowMap m = row_map(a.r, NTHR, threadIdx.x);
    if (!m.active) return;
    ChanCtx c;
    load_ctx<T, KIND>(a, n, m.cv, c);
    float cq[8], cp[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int g = (m.cv * 8 + k) / gs;
        cq[k] = a.mstat[(n * kGroups + g) * 2];
        cp[k] = a.mstat[(n * kGroups + g) * 2 + 1];
    }
    const int uy_end = min(UH, (int)(blockIdx.x + 1) * rows_pb);
    constexpr int UB = P == 4 ? 1 : 2;
    for (int uy = blockIdx.x * rows_pb; uy < uy_end; ++uy) {
        for (int ux0 = m.x0; ux0 < UW; ux0 += UB * m.xstep) {
            float rv[UB][P][8], dz[UB][P][8], act[8], dl;
#pragma unroll
            for (int u = 0; u < UB; ++u) {
                const int ux = min(ux0 + u * m.xstep, UW - 1);
                unit_grads<T, KIND, P>(a, c, n, uy, ux, m.cv, rv[u], dz[u], dl, act);
            }
#pragma unroll
            for (int u = 0; u < UB; ++u) {
                const int ux = ux0 + u * m.xstep;
                if (ux >= UW) continue;
#pragma unroll
                for (int q = 0; q < P; ++q) {
                    const int y = P == 4 ? 2 * uy + (q >> 1) : uy, x = P == 4 ? 2 * ux + (q & 1) : ux;
                    float o[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) o[k] = fmaf(c.sc[k], dz[u][q][k], fmaf(cp[k], rv[u][q][k], cq[k]));
                    store8(a.dr.at(n, y, x, m.cv * 8), o);
                }
            }
        }
    }
}

template <typename T, int KIND>
int run(const GnBwdArgs<T>& a, cudaStream_t s) {
    constexpr int P = KindTraits<KIND>::pool ? 4 : 1;
    const int C = a.r.C, nv = C / 8;
    if (nv > NTHR) { set_error("gn_bwd: C = %d is too wide", C); return -1; }
    const int UH = a.r.H / (P == 4 ? 2 : 1), UW = a.r.W / (P == 4 ? 2 : 1);
    const int rows_pb = rows_per_block(UW, nv, NTHR, UNITS_PER_THREAD);
    const int slots = ceil_div(UH, rows_pb);
    const int pcs = KIND == SRC_BLEND_PS ? C / 4 : C;
    const double inv_count = 1.0 / ((double)(C / kGroups) * a.r.H * a.r.W);
    const int rows_ap = rows_per_block(UW, nv, NTHR, 8);
    // Samples are processed in chunks small enough for the r / g tensors of a chunk to stay in L2 between the reduce pass
    // and the apply pass (GroupNorm couples a whole sample, so the two passes cannot be one): 5 -> 3 HBM passes.
    const double sample_bytes = 3.0 * sizeof(T) * (double)C * a.r.H * a.r.W;          // r + g read, dr written
    const double chunk_d = gn_bwd_chunk_bytes() / sample_bytes;
    int chunk = chunk_d >= (double)a.r.N ? a.r.N : (int)chunk_d;
    if (chunk < 1) chunk = 1;
    for (int n0 = 0; n0 < a.r.N; n0 += chunk) {
        const int nn = a.r.N - n0 < chunk ? a.r.N - n0 : chunk;
        gn_bwd_reduce_kernel<T, KIND><<<dim3(slots, nn), NTHR, 0, s>>>(a, rows_pb, n0);
        MRISR_LAUNCH_CHECK();
        gn_bwd_finalize_kernel<T, KIND><<<nn * kGroups, FTHR, 0, s>>>(a, slots, inv_count, n0, nn, 0);
        MRISR_LAUNCH_CHECK();
        gn_bwd_apply_kernel<T, KIND><<<dim3(ceil_div(UH, rows_ap), nn), NTHR, 0, s>>>(a, rows_ap, n0);
        MRISR_LAUNCH_CHECK();
    }
    gn_bwd_finalize_kernel<T, KIND><<<pcs, FTHR, 0, s>>>(a, slots, inv_count, 0, a.r.N, 1);
    MRISR_LAUNCH_CHECK();
    return 0;
}
}  // namespace

int gn_bwd_slots(int H, int W, int C) {
    // upper bound over both unit shapes (pool kinds use half the rows and half the columns)
    const int full = ceil_div(H, rows_per_block(W, C / 8, NTHR, UNITS_PER_THREAD));
    const int pool = ceil_div(H / 2, rows_per_block(W / 2, C / 8, NTHR, UNITS_PER_THREAD));
    return full > pool ? full : pool;
}

template <typename T>
int launch_gn_bwd(const GnBwdArgs<T>& a, cudaStream_t s) {
    const char* tag = "gn_bwd";
    char tbuf[96];
    if (prof_on()) {
        snprintf(tbuf, sizeof(tbuf), "gn_bwd:kind%d:%d@%dx%d", a.kind, a.r.C, a.r.H, a.r.W);
        tag = prof_intern(tbuf);
    }
    ProfScope prof_(tag, s);
    switch (a.kind) {
        case SRC_DIRECT: return run<T, SRC_DIRECT>(a, s);
        case SRC_DIRECT_POOL: return run<T, SRC_DIRECT_POOL>(a, s);
        case SRC_OUTER: return run<T, SRC_OUTER>(a, s);
        case SRC_BLEND_BIL: return run<T, SRC_BLEND_BIL>(a, s);
        case SRC_BLEND_PS: return run<T, SRC_BLEND_PS>(a, s);
        default: set_error("gn_bwd: bad source kind %d", a.kind); return -1;
    }
}
template int launch_gn_bwd<float>(const GnBwdArgs<float>&, cudaStream_t);
template int launch_gn_bwd<__nv_bfloat16>(const GnBwdArgs<__nv_bfloat16>&, cudaStream_t);

}  // namespace mrisr
