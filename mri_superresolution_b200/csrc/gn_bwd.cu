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
#include <algorithm>
#include "kernels.cuh"
#include "gn_bwd_trk.h"

namespace mrisr {
namespace {
constexpr int NTHR = 256;
constexpr int UNITS_PER_THREAD = 32;
constexpr int FTHR = 256;         // threads of the finalize kernel

template <int KIND> struct KindTraits { static constexpr bool pool = (KIND == SRC_DIRECT_POOL || KIND == SRC_POOL_ONLY); };

// The blend kinds also accumulate  sum g * lrelu(z)  per channel (third partial-sum row, as SRC_OUTER does for the 1x1
// weights) and the finalize kernel writes the per-parameter-channel totals to `dwvec`; the caller (unet_run.cu) forms
// d(alpha) = s(1-s) * (sum over the bilinear branch - sum over the pixel-shuffle branch) from the two vectors.  Measured on
// B200 (round 2, profiles/r02_staged_variants.md): the separate blend_dalpha pass it replaces cost 0.363 ms of the 17.6 ms
// step (it re-read both branches and the gradient for one scalar); the extra sums cost 0.03 ms in the two blend launches.
constexpr bool kDalpha = true;
template <int KIND> struct ThirdRow {
    static constexpr bool blend = kDalpha && (KIND == SRC_BLEND_BIL || KIND == SRC_BLEND_PS);
    static constexpr bool on = KIND == SRC_OUTER || blend;
};


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
            if (ThirdRow<KIND>::blend) {
#pragma unroll
                for (int k = 0; k < 8; ++k) act_out[k] = dA[k] * lrelu(z[q][k]);
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) dA[k] *= al;
        } else if (KIND == SRC_BLEND_PS) {
            const float om = 1.f - 1.f / (1.f + __expf(-a.alpha[0]));
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int ch = cv * 8 + k, cs = ch >> 2, i = (ch >> 1) & 1, j = ch & 1;
                const float graw = to_f(*a.g.at(n, 2 * y + i, 2 * x + j, cs));
                act_out[k] = graw * lrelu(z[q][k]);
                dA[k] = om * graw;
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
                    } else if (ThirdRow<KIND>::blend) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) s3[k] += act[u][k];
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
            if (ThirdRow<KIND>::on) d3 += (double)q[2 * row + ch];
            if (KIND == SRC_OUTER && pc == 0) d4 += (double)q[C];
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
            } else if (ThirdRow<KIND>::blend) {
                if (a.dwvec) a.dwvec[pc] = (float)d3;            // sum g * lrelu(z) of this parameter channel (d(alpha) term)
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

// ------------------------------------------------------------------------------------------------------------------
// Fused form (bf16): ONE kernel does reduce -> group constants -> apply.  GroupNorm couples all pixels of a (sample, group),
// so the unit of independence is a sample (NHWC) or one 8-channel plane of a sample (planar-8, C <= 64: groups never straddle
// planes).  A thread-block CLUSTER owns a unit: every CTA reduces its share of the rows, the per-channel partial sums are
// exchanged through distributed shared memory (fixed rank order: deterministic), every CTA derives (q, p) of its groups and
// applies dr = sc*dz + p*r + q to the same rows.
// Measured on B200 (batch 64, profiles/r01_gn_bwd_fused.md): clusters of 2 are fastest (gn_bwd 6.8 -> 5.0 ms per step); larger
// clusters lose more to gang scheduling than their bigger shared-memory share of a unit wins back, and with 2 CTAs the
// operands kept in shared memory between the phases (K slots) are a few percent of a unit, so the second phase still
// re-reads most operands (5 passes, now at ~70% of the HBM copy rate instead of 45%).  A persistent "streamed" form that sizes the
// units in flight to the L2 (groups of CTAs per unit, arrival counters instead of cluster barriers, the next unit's reduce between
// arrive and wait) was built and measured in round 2 and LOST (gn_bwd 4.6 -> 7.2 ms per step): the L2 budget forces ~110 KB
// phases per CTA, whose fixed per-phase instructions made it issue-bound (66 % issue-active, L2 hit rate 30 %, DRAM traffic still
// 4.2 passes); deleted, numbers in profiles/r02_gn_bwd_stream.md.
constexpr int kFtMax = 512;             // threads per CTA of the fused kernel: 512 (2 CTAs / SM) or 256 (4 CTAs / SM)
constexpr int kMaxUnitCh = 512;         // channels of one unit

template <int P, bool POOL> struct Raw { uint4 r[P]; uint4 a[P]; uint4 gp; };

__device__ __forceinline__ void unpack8(const uint4& r, float (&v)[8]) {
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        v[2 * i] = __uint_as_float(w[i] << 16);
        v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
}

// same arithmetic as unit_grads, operands taken from the raw words
template <int KIND, int P>
__device__ __forceinline__ void grads_from_raw(const ChanCtx& c, const Raw<P, KindTraits<KIND>::pool>& w, float al, const float (&wv)[8],
                                               float (&rv)[P][8], float (&dz)[P][8], float& dl_out, float (&act_out)[8]) {
    using T = __nv_bfloat16;
    float z[P][8];
#pragma unroll
    for (int q = 0; q < P; ++q) {
        unpack8(w.r[q], rv[q]);
#pragma unroll
        for (int k = 0; k < 8; ++k) z[q][k] = fmaf(rv[q][k], c.sc[k], c.sh[k]);
        float dA[8];
        if (KIND == SRC_DIRECT || KIND == SRC_DIRECT_POOL) {
            unpack8(w.a[q], dA);
        } else if (KIND == SRC_BLEND_BIL) {
            unpack8(w.a[q], dA);
            if (ThirdRow<KIND>::blend) {
#pragma unroll
                for (int k = 0; k < 8; ++k) act_out[k] = dA[k] * lrelu(z[q][k]);
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) dA[k] *= al;
        } else if (KIND == SRC_BLEND_PS) {
            const uint32_t t[4] = {w.a[q].x, w.a[q].y, w.a[q].z, w.a[q].w};
#pragma unroll
            for (int ij = 0; ij < 4; ++ij) {
                const float g0 = __uint_as_float(t[ij] << 16), g1 = __uint_as_float(t[ij] & 0xffff0000u);
                act_out[ij] = g0 * lrelu(z[q][ij]);
                act_out[4 + ij] = g1 * lrelu(z[q][4 + ij]);
                dA[ij] = al * g0;
                dA[4 + ij] = al * g1;
            }
        } else if (KIND == SRC_OUTER) {
            const float yy = __uint_as_float(w.a[q].x);
            const float dl = __uint_as_float(w.a[q].y) * yy * (1.f - yy);
            dl_out = dl;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                dA[k] = dl * wv[k];
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
        float gp[8];
        unpack8(w.gp, gp);
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

struct FusedParams {
    int cs;            // CTAs per cluster (= partial-sum slots per sample)
    int ncv;           // channel vectors per unit (1: one plane of a planar tensor; C/8: a whole NHWC sample)
    int ups;           // units per sample
    int rows_pc;       // unit rows per CTA
    int slots;         // shared-memory slots per thread (K, or K + S)
    int K;             // units per thread kept in shared memory between the phases (plus S streaming slots)
    double inv_count;
    // Incremental form (the plain-tensor kinds, kIncr below): operand addresses by constant steps (gn_bwd_trk.h) instead of a
    // division and 64-bit index arithmetic per unit (~100 of the 203 SASS instructions of the kind-0 phase-1 loop; the kernel
    // was issue-bound at 417 thread instructions per 16-byte unit, profiles/r01_gn_bwd_issue_bound.md).
    TrkStep ts;
    int psr, psg, psd, psh;      // pixel strides (elements) of r, g, dr, gpool
};

// Which kinds use the incremental addressing: measured per kind on B200 (round 2, batch 64 train step, ms per step summed over
// the launches of the kind, division form -> incremental form): plain tensor 2.39 -> 1.90, + max-pool routing 0.748 -> 0.746,
// final-head outer product 0.639 -> 0.710, blend 0.571 -> 0.490 and pixel-shuffle 0.706 -> 0.625 -- but the last two lose
// 0.39 ms once they also carry the d(alpha) sums (register pressure at 64 registers per thread), which is more than the
// 0.16 ms the incremental form wins there and more than the blend_dalpha pass the sums replace, so they keep the division form.
template <int KIND> struct Incr { static constexpr bool on = KIND == SRC_DIRECT || KIND == SRC_DIRECT_POOL; };

__device__ __forceinline__ void unit_row_col(int pi, int UW, int& row, int& col) {
    row = pi / UW;
    col = pi - row * UW;
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
// shared-state-space forms (32-bit shared address) for the incremental build: no generic -> shared conversion per access
[[maybe_unused]] __device__ __forceinline__ void cp_async16_s(uint32_t sdst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sdst), "l"(gsrc) : "memory");
}
[[maybe_unused]] __device__ __forceinline__ void cp_async4_s(uint32_t sdst, const void* gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sdst), "l"(gsrc) : "memory");
}
[[maybe_unused]] __device__ __forceinline__ uint4 ld_shared_v4(uint32_t saddr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr) : "memory");
    return v;
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Every operand goes global -> shared memory with cp.async into thread-private slots (no registers held while in flight):
// K "cached" slots keep the first K units of the thread for the second phase, S "streaming" slots form a depth-S software
// pipeline for the rest.  The second phase walks the streamed units in reverse order (most recently read first: L2), and its
// first S reloads are issued before the cluster barriers so the exchange latency is hidden behind them.
template <int KIND, int FT, int S>
__global__ void __launch_bounds__(FT, (KindTraits<KIND>::pool ? 1 : 2) * (512 / FT)) gn_bwd_fused_kernel(GnBwdArgs<__nv_bfloat16> a, FusedParams fp) {
    using T = __nv_bfloat16;
    constexpr bool POOL = KindTraits<KIND>::pool;
    constexpr bool kIncr = Incr<KIND>::on;
    constexpr int P = POOL ? 4 : 1;
    constexpr int RAWV = 2 * P + (POOL ? 1 : 0);
    constexpr int NW = FT / 32;
    extern __shared__ uint4 dyn_smem[];
    __shared__ float s_ch[3][kMaxUnitCh];
    __shared__ double s_tot[2][kMaxUnitCh];
    __shared__ float s_q[kGroups], s_p[kGroups], s_wdl[NW];
    static_assert(FT <= kFtMax, "");

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    uint32_t rank, unit;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(unit));
    const int ncv = fp.ncv, nchl = ncv * 8;
    const int n = unit / fp.ups, cv0 = (unit % fp.ups) * ncv;
    const int C = a.r.C, gs = C / kGroups;
    const int UW = P == 4 ? a.r.W / 2 : a.r.W, UH = P == 4 ? a.r.H / 2 : a.r.H;
    const int r0 = min(UH, (int)rank * fp.rows_pc), r1 = min(UH, r0 + fp.rows_pc);
    const int npx = (r1 - r0) * UW;                    // units of this CTA per channel vector
    const int cvl = tid % ncv, cv = cv0 + cvl;
    const int p0 = tid / ncv, pstep = FT / ncv;
    const int nu = p0 < npx ? (npx - p0 + pstep - 1) / pstep : 0;      // units of this thread
    const int K = fp.K;
    uint4* cache = dyn_smem;                           // [K + S][RAWV][FT]
    float* s_w = reinterpret_cast<float*>(dyn_smem + (size_t)fp.slots * RAWV * FT);      // [NW][3][nchl]

    ChanCtx c;
    load_ctx<T, KIND>(a, n, cv, c);
    float al = 0.f, wv[8];
    if (KIND == SRC_BLEND_BIL) al = 1.f / (1.f + __expf(-a.alpha[0]));
    if (KIND == SRC_BLEND_PS) al = 1.f - 1.f / (1.f + __expf(-a.alpha[0]));
#pragma unroll
    for (int k = 0; k < 8; ++k) wv[k] = KIND == SRC_OUTER ? a.wvec[cv * 8 + k] : 0.f;

    // incremental form (kIncr): per-thread base pointers at (first row of this CTA, column 0) and the pixel strides in elements
    [[maybe_unused]] constexpr int PM = P == 4 ? 2 : 1;                 // pixels per unit along each axis
    [[maybe_unused]] const int pitch = a.r.W + 2, hpitch = UW + 2, gpitch = a.g.W + 2;
    [[maybe_unused]] const int psr = fp.psr, psg = fp.psg, psd = fp.psd, psh = fp.psh;      // pixel strides in elements
    // slot handles: the slot index (default build) or the shared-memory byte address of this thread's first vector in the slot
    // (kIncr: advanced by a constant with one wrap over the S streaming slots instead of slot_of(i) and a pointer product)
    [[maybe_unused]] constexpr uint32_t kVec = FT * 16, kSlot = RAWV * FT * 16;
    [[maybe_unused]] const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(cache) + tid * 16;
    [[maybe_unused]] const uint32_t swrap0 = sbase + (uint32_t)K * kSlot, swrap1 = swrap0 + S * kSlot;
    auto hnext = [&](uint32_t h) { h += kSlot; return h == swrap1 ? swrap0 : h; };
    [[maybe_unused]] const T* rb = nullptr; [[maybe_unused]] const T* gb = nullptr; [[maybe_unused]] const T* hb = nullptr; [[maybe_unused]] T* db = nullptr;
    if constexpr (kIncr) {
        rb = a.r.at(n, PM * r0, 0, cv * 8);
        db = a.dr.at(n, PM * r0, 0, cv * 8);
        if (KIND == SRC_DIRECT || KIND == SRC_DIRECT_POOL) gb = a.g.at(n, PM * r0, 0, a.coff + cv * 8);
        if (KIND == SRC_BLEND_BIL) gb = a.g.at(n, r0, 0, cv * 8);
        if (KIND == SRC_BLEND_PS) gb = a.g.at(n, 2 * r0, 0, 2 * cv);
        if (POOL) hb = a.gpool.at(n, r0, 0, cv * 8);
    }
    auto issue = [&](int i, uint32_t h, const UnitTrk& t) {              // unit i of this thread (at position t) -> slot handle h
        if constexpr (kIncr) {
#pragma unroll
            for (int q = 0; q < P; ++q) {
                const int po = PM * t.px + (q >> 1) * pitch + (q & 1);         // pixels from the base, r / g / dr geometry
                cp_async16_s(h + q * kVec, rb + po * psr);
                const uint32_t sa = h + (P + q) * kVec;
                if (KIND == SRC_DIRECT || KIND == SRC_DIRECT_POOL || KIND == SRC_BLEND_BIL) {
                    cp_async16_s(sa, gb + po * psg);
                } else if (KIND == SRC_BLEND_PS) {
#pragma unroll
                    for (int ij = 0; ij < 4; ++ij)
                        cp_async4_s(sa + 4 * ij, gb + (2 * t.gpx + (ij >> 1) * gpitch + (ij & 1)) * psg);
                } else if (KIND == SRC_OUTER) {
                    const int64_t pix = ((int64_t)n * a.r.H + r0) * a.r.W + (p0 + i * pstep);
                    cp_async4_s(sa, a.y + pix);
                    cp_async4_s(sa + 4, a.dy + pix);
                }
            }
            if (POOL) cp_async16_s(h + 2 * P * kVec, hb + t.hpx * psh);
            return;
        }
        const int sidx = (int)h;
        uint4* sl = cache + (size_t)sidx * RAWV * FT + tid;
        const int pi = p0 + i * pstep;
        int uy, ux;
        unit_row_col(pi, UW, uy, ux);
        uy += r0;
#pragma unroll
        for (int q = 0; q < P; ++q) {
            const int y = P == 4 ? 2 * uy + (q >> 1) : uy, x = P == 4 ? 2 * ux + (q & 1) : ux;
            cp_async16(sl + q * FT, a.r.at(n, y, x, cv * 8));
            uint4* sa = sl + (P + q) * FT;
            if (KIND == SRC_DIRECT || KIND == SRC_DIRECT_POOL) {
                cp_async16(sa, a.g.at(n, y, x, a.coff + cv * 8));
            } else if (KIND == SRC_BLEND_BIL) {
                cp_async16(sa, a.g.at(n, y, x, cv * 8));
            } else if (KIND == SRC_BLEND_PS) {
                // channels 8cv..8cv+7 of the unshuffled tensor = channels (2cv, 2cv+1) of g at the four sub-pixels (i, j)
#pragma unroll
                for (int ij = 0; ij < 4; ++ij)
                    cp_async4(reinterpret_cast<uint32_t*>(sa) + ij, a.g.at(n, 2 * y + (ij >> 1), 2 * x + (ij & 1), 2 * cv));
            } else if (KIND == SRC_OUTER) {
                const int64_t pix = ((int64_t)n * a.r.H + y) * a.r.W + x;
                cp_async4(reinterpret_cast<uint32_t*>(sa), a.y + pix);
                cp_async4(reinterpret_cast<uint32_t*>(sa) + 1, a.dy + pix);
            }
        }
        if (POOL) cp_async16(sl + 2 * P * FT, a.gpool.at(n, uy, ux, cv * 8));
    };
    auto fetch = [&](uint32_t h, Raw<P, POOL>& w) {
        if constexpr (kIncr) {
#pragma unroll
            for (int q = 0; q < P; ++q) { w.r[q] = ld_shared_v4(h + q * kVec); w.a[q] = ld_shared_v4(h + (P + q) * kVec); }
            if (POOL) w.gp = ld_shared_v4(h + 2 * P * kVec);
            return;
        }
        const int sidx = (int)h;
        const uint4* sl = cache + (size_t)sidx * RAWV * FT + tid;
#pragma unroll
        for (int q = 0; q < P; ++q) { w.r[q] = sl[q * FT]; w.a[q] = sl[(P + q) * FT]; }
        if (POOL) w.gp = sl[2 * P * FT];
    };
    auto slot_of = [&](int i) { return i < K ? i : K + (i - K) % S; };

    // ---- phase 1: partial sums of this CTA's rows
    float s1[8], s2[8], s3[8], sdl = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s1[k] = s2[k] = s3[k] = 0.f;
    UnitTrk ti{0, 0, 0, 0};                            // position of the next unit to issue (kIncr)
    if constexpr (kIncr) ti = trk_at(p0, UW, pitch, hpitch, gpitch);
    [[maybe_unused]] uint32_t hi = sbase, hf = sbase;  // handles of the next slot to issue into / to fetch from (kIncr)
#pragma unroll
    for (int i = 0; i < S; ++i) {
        if (i < nu) {
            issue(i, kIncr ? hi : (uint32_t)slot_of(i), ti);
            if constexpr (kIncr) { trk_fwd(ti, fp.ts, UW); hi = hnext(hi); }
        }
        cp_async_commit();
    }
    for (int i = 0; i < nu; ++i) {
        cp_async_wait<S - 1>();
        Raw<P, POOL> w;
        fetch(kIncr ? hf : (uint32_t)slot_of(i), w);
        if constexpr (kIncr) hf = hnext(hf);
        float rv[P][8], dz[P][8], act[8], dl = 0.f;
        grads_from_raw<KIND, P>(c, w, al, wv, rv, dz, dl, act);
#pragma unroll
        for (int q = 0; q < P; ++q)
#pragma unroll
            for (int k = 0; k < 8; ++k) { s1[k] += dz[q][k]; s2[k] = fmaf(dz[q][k], rv[q][k], s2[k]); }
        if (KIND == SRC_OUTER) {
#pragma unroll
            for (int k = 0; k < 8; ++k) s3[k] += dl * act[k];
            if (cvl == 0) sdl += dl;
        } else if (ThirdRow<KIND>::blend) {
#pragma unroll
            for (int k = 0; k < 8; ++k) s3[k] += act[k];
        }
        if (i + S < nu) {                                      // the streaming slot just consumed (or a fresh one)
            issue(i + S, kIncr ? hi : (uint32_t)slot_of(i + S), ti);
            if constexpr (kIncr) { trk_fwd(ti, fp.ts, UW); hi = hnext(hi); }
        }
        cp_async_commit();
    }
    // reloads of the streamed units for the second phase, newest first: unit nu-1-t -> streaming slot t % S
    const int nstream = nu > K ? nu - K : 0;
    UnitTrk ta = ti;                                   // every unit was issued once: ti (and ta) stand at unit nu
#pragma unroll
    for (int t = 0; t < S; ++t) {
        if (t < nstream) {
            if constexpr (kIncr) trk_bwd(ti, fp.ts, UW);
            issue(nu - 1 - t, kIncr ? swrap0 + t * kSlot : (uint32_t)(K + t), ti);
        }
        cp_async_commit();
    }
    // lanes with equal (lane % ncv) share a channel vector: butterfly over the other lane bits, then across warps in order
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        for (int o = 16; o >= ncv; o >>= 1) {
            s1[k] += __shfl_xor_sync(0xffffffffu, s1[k], o);
            s2[k] += __shfl_xor_sync(0xffffffffu, s2[k], o);
            if (ThirdRow<KIND>::on) s3[k] += __shfl_xor_sync(0xffffffffu, s3[k], o);
        }
    }
    if (lane < ncv) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            s_w[(wid * 3 + 0) * nchl + lane * 8 + k] = s1[k];
            s_w[(wid * 3 + 1) * nchl + lane * 8 + k] = s2[k];
            s_w[(wid * 3 + 2) * nchl + lane * 8 + k] = s3[k];
        }
    }
    if (KIND == SRC_OUTER) {
        sdl = warp_sum(sdl);
        if (lane == 0) s_wdl[wid] = sdl;
    }
    __syncthreads();
    const int64_t row = (int64_t)(C + 8);
    float* out = a.part + ((int64_t)n * fp.cs + rank) * 3 * row;
    // with ncv < 32 a warp covers channel vector (lane % ncv) only when ncv divides 32; warps hold every local vector
    for (int i = tid; i < 3 * nchl; i += FT) {
        const int which = i / nchl, chl = i % nchl;
        float t = 0.f;
        if (ncv <= 32) {
            for (int w2 = 0; w2 < NW; ++w2) t += s_w[(w2 * 3 + which) * nchl + chl];
        } else {
            // ncv = 64: lane l of warp w holds vector (w * 32 + l) % 64 -> warps of equal parity share vectors
            const int v = chl >> 3, par = v >> 5;
            for (int w2 = par; w2 < NW; w2 += 2) t += s_w[(w2 * 3 + which) * nchl + ((v & 31) * 8 + (chl & 7))];
        }
        s_ch[which][chl] = t;
        out[which * row + cv0 * 8 + chl] = t;
    }
    if (KIND == SRC_OUTER && tid == 0) {
        float t = 0.f;
        for (int w2 = 0; w2 < NW; ++w2) t += s_wdl[w2];
        if (cv0 == 0) out[C] = t;
    }
    // ---- exchange: per-channel totals of the unit in rank order, then the constants of the unit's groups
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    for (int i = tid; i < 2 * nchl; i += FT) {
        const int which = i / nchl, chl = i % nchl;
        const uint32_t local = (uint32_t)__cvta_generic_to_shared(&s_ch[which][chl]);
        double t = 0.0;
        for (int rk = 0; rk < fp.cs; ++rk) {
            uint32_t remote;
            float v;
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(rk));
            asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote) : "memory");
            t += (double)v;
        }
        s_tot[which][chl] = t;
    }
    __syncthreads();
    const int ngl = nchl >= gs ? nchl / gs : 0;        // whole groups inside the unit (host guarantees >= 1)
    if (tid < ngl) {
        const int g = (cv0 * 8) / gs + tid, ng = n * kGroups + g;
        const double mean = (double)a.stats[ng * 2], rstd = (double)a.stats[ng * 2 + 1];
        double m1 = 0.0, m2 = 0.0;
        for (int j = 0; j < gs; ++j) {
            const int chl = tid * gs + j, ch = cv0 * 8 + chl;
            const double gm = (double)a.gamma[KIND == SRC_BLEND_PS ? ch / 4 : ch];
            m1 += gm * s_tot[0][chl];
            m2 += gm * (s_tot[1][chl] - mean * s_tot[0][chl]);
        }
        m1 *= fp.inv_count;
        m2 *= rstd * fp.inv_count;
        s_q[tid] = (float)(-rstd * m1 + mean * rstd * rstd * m2);
        s_p[tid] = (float)(-rstd * rstd * m2);
    }
    // second cluster barrier: nobody's shared memory is read remotely after it (CTAs may exit), and s_q / s_p are visible
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");

    // ---- phase 2: dr = sc*dz + p*r + q over the same rows (cached operands first, then the streamed ones)
    float cq[8], cp[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int gl = (cvl * 8 + k) / gs;
        cq[k] = s_q[gl];
        cp[k] = s_p[gl];
    }
    auto apply = [&](int i, const Raw<P, POOL>& w, const UnitTrk& t) {
        [[maybe_unused]] int uy = 0, ux = 0;
        if constexpr (!kIncr) {
            const int pi = p0 + i * pstep;
            unit_row_col(pi, UW, uy, ux);
            uy += r0;
        }
        float rv[P][8], dz[P][8], act[8], dl = 0.f;
        grads_from_raw<KIND, P>(c, w, al, wv, rv, dz, dl, act);
#pragma unroll
        for (int q = 0; q < P; ++q) {
            float o[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = fmaf(c.sc[k], dz[q][k], fmaf(cp[k], rv[q][k], cq[k]));
            if constexpr (kIncr) {
                store8(db + (PM * t.px + (q >> 1) * pitch + (q & 1)) * psd, o);
            } else {
                const int y = P == 4 ? 2 * uy + (q >> 1) : uy, x = P == 4 ? 2 * ux + (q & 1) : ux;
                store8(a.dr.at(n, y, x, cv * 8), o);
            }
        }
    };
    UnitTrk tc{0, 0, 0, 0};                            // cached units: ascending from the thread's first unit
    if constexpr (kIncr) tc = trk_at(p0, UW, pitch, hpitch, gpitch);
    hf = sbase;
    for (int i = 0; i < K && i < nu; ++i) {
        Raw<P, POOL> w;
        fetch(kIncr ? hf : (uint32_t)i, w);
        apply(i, w, tc);
        if constexpr (kIncr) { trk_fwd(tc, fp.ts, UW); hf += kSlot; }
    }
    hf = swrap0;
    for (int t = 0; t < nstream; ++t) {                // streamed units: descending from unit nu - 1
        cp_async_wait<S - 1>();
        Raw<P, POOL> w;
        const uint32_t hs = kIncr ? hf : (uint32_t)(K + t % S);       // streaming slot t % S: fetched, then refilled
        fetch(hs, w);
        if constexpr (kIncr) { trk_bwd(ta, fp.ts, UW); hf = hnext(hf); }
        apply(nu - 1 - t, w, ta);
        if (t + S < nstream) {
            if constexpr (kIncr) trk_bwd(ti, fp.ts, UW);
            issue(nu - 1 - (t + S), hs, ti);
        }
        cp_async_commit();
    }
}

struct FusedPlan { bool ok; FusedParams fp; size_t smem; int units; int ft; int depth; };

inline int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}

inline int fused_threads() {
    return 512;                       // 2 CTAs / SM; 256 (4 CTAs / SM) measured slower in round 1 (profiles/r01_gn_bwd_fused.md)
}

template <int KIND>
FusedPlan fused_plan(const GnBwdArgs<__nv_bfloat16>& a) {
    const int FT = fused_threads();
    constexpr bool POOL = KindTraits<KIND>::pool;
    constexpr int P = POOL ? 4 : 1, RAWV = 2 * P + (POOL ? 1 : 0);
    FusedPlan pl;
    pl.ok = false;
    static const int enabled = env_int("MRISR_GNBWD_FUSED", 1);
    // two CTAs per SM (64 registers) for the one-pixel kinds; the 2x2-window kinds need 128 registers -> one CTA with all the SM's
    // shared memory
    const int smem_kb = FT == 512 ? (POOL ? 190 : 96) : (POOL ? 95 : 40);
    if (!enabled) return pl;
    const int C = a.r.C, nv = C / 8, gs = C / kGroups;
    const int ncv = a.r.planar() ? 1 : nv;
    if (C % 8 || ncv * 8 > kMaxUnitCh || FT % ncv || (ncv * 8) % gs) return pl;          // groups must not straddle units
    const int UH = a.r.H / (P == 4 ? 2 : 1), UW = a.r.W / (P == 4 ? 2 : 1);
    if (UH < 1 || UW < 1) return pl;
    const size_t red_bytes = (size_t)(FT / 32) * 3 * ncv * 8 * sizeof(float);
    const size_t budget = (size_t)smem_kb * 1024 > red_bytes + 4096 ? (size_t)smem_kb * 1024 - red_bytes : 4096;
    const int kmax = (int)(budget / ((size_t)FT * RAWV * 16));
    int cs = 2;                          // clusters of 2 won the sweep in both rounds (4: +0.11 ms, 8: +0.42 ms per step under graph replay)
    while (cs > UH && cs > 1) cs >>= 1;
    if (cs != 1 && cs != 2 && cs != 4 && cs != 8 && cs != 16) return pl;
    pl.ft = FT;
    pl.fp.cs = cs;
    pl.fp.ncv = ncv;
    pl.fp.ups = nv / ncv;
    pl.fp.rows_pc = ceil_div(UH, cs);
    const int iters = (int)ceil_div64((int64_t)pl.fp.rows_pc * UW * ncv, FT);
    // pipeline depth S (= streaming slots): 2 for the 2x2-window kinds (9 vectors per unit), else 4
    const int S = POOL ? 2 : 4;
    const int k_env = -1;
    int slots;                                                     // shared-memory slots per thread
    if (iters <= kmax && k_env < 0) { pl.fp.K = iters; slots = iters; }      // everything stays resident: no streaming slots
    else {
        if (kmax < S) return pl;
        pl.fp.K = iters < kmax - S ? iters : kmax - S;
        if (k_env >= 0 && k_env < pl.fp.K) pl.fp.K = k_env;
        slots = pl.fp.K + S;
    }
    pl.depth = S;
    pl.fp.inv_count = 1.0 / ((double)gs * a.r.H * a.r.W);
    pl.fp.ts = trk_steps(FT / ncv, UW, a.r.W + 2, a.g.W + 2);
    pl.fp.psr = a.r.ps ? a.r.ps : a.r.C;
    pl.fp.psg = a.g.ps ? a.g.ps : a.g.C;
    pl.fp.psd = a.dr.ps ? a.dr.ps : a.dr.C;
    pl.fp.psh = POOL ? (a.gpool.ps ? a.gpool.ps : a.gpool.C) : 0;
    if (Incr<KIND>::on) {
        // 32-bit element offsets inside a sample, and the operand geometries the incremental form assumes
        auto elems = [](const Act<__nv_bfloat16>& t) { return (double)t.sample_pixels() * (t.ps ? t.ps : t.C); };
        const bool same_res = KIND == SRC_DIRECT || KIND == SRC_DIRECT_POOL || KIND == SRC_BLEND_BIL;
        if (elems(a.r) >= 2147483648.0 || elems(a.dr) >= 2147483648.0) return pl;
        if (a.dr.W != a.r.W || a.dr.H != a.r.H) return pl;
        if (KIND != SRC_OUTER && elems(a.g) >= 2147483648.0) return pl;
        if (same_res && (a.g.W != a.r.W || a.g.H != a.r.H)) return pl;
        if (KIND == SRC_BLEND_PS && (a.g.W != 2 * a.r.W || a.g.H != 2 * a.r.H)) return pl;
        if (POOL && (a.gpool.W != UW || a.gpool.H != UH || elems(a.gpool) >= 2147483648.0)) return pl;
    }
    pl.smem = (size_t)slots * RAWV * FT * 16 + red_bytes;
    pl.fp.slots = slots;
    pl.units = a.r.N * pl.fp.ups;
    pl.ok = true;
    return pl;
}

template <int KIND, int FT, int S>
int run_fused(const GnBwdArgs<__nv_bfloat16>& a, const FusedPlan& pl, cudaStream_t s) {
    static PerDeviceFlag attr_done_pd; bool& attr_done = attr_done_pd.cur();
    if (!attr_done) {
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(gn_bwd_fused_kernel<KIND, FT, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(gn_bwd_fused_kernel<KIND, FT, S>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        attr_done = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(pl.units * pl.fp.cs));
    cfg.blockDim = dim3(FT);
    cfg.dynamicSmemBytes = pl.smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = pl.fp.cs;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    MRISR_CUDA_CHECK(cudaLaunchKernelEx(&cfg, gn_bwd_fused_kernel<KIND, FT, S>, a, pl.fp));
    MRISR_LAUNCH_CHECK();
    const int pcs = KIND == SRC_BLEND_PS ? a.r.C / 4 : a.r.C;
    gn_bwd_finalize_kernel<__nv_bfloat16, KIND><<<pcs, FTHR, 0, s>>>(a, pl.fp.cs, pl.fp.inv_count, 0, a.r.N, 1);
    MRISR_LAUNCH_CHECK();
    return 0;
}

template <typename T, int KIND>
int run(const GnBwdArgs<T>& a, cudaStream_t s) {
    constexpr int P = KindTraits<KIND>::pool ? 4 : 1;
    const int C = a.r.C, nv = C / 8;
    if (nv > NTHR) { set_error("gn_bwd: C = %d is too wide", C); return -1; }
    if constexpr (sizeof(T) == 2) {
        const FusedPlan fpl = fused_plan<KIND>(a);
        if (fpl.ok) {
            constexpr int S0 = KindTraits<KIND>::pool ? 2 : 4;
            return run_fused<KIND, 512, S0>(a, fpl, s);
        }
    }
    const int UH = a.r.H / (P == 4 ? 2 : 1), UW = a.r.W / (P == 4 ? 2 : 1);
    const int rows_pb = rows_per_block(UW, nv, NTHR, UNITS_PER_THREAD);
    const int slots = ceil_div(UH, rows_pb);
    const int pcs = KIND == SRC_BLEND_PS ? C / 4 : C;
    const double inv_count = 1.0 / ((double)(C / kGroups) * a.r.H * a.r.W);
    const int rows_ap = rows_per_block(UW, nv, NTHR, 8);
    // (per-chunk reduce + apply for L2 reuse was measured and rejected: profiles/r01_gn_bwd_chunking.md)
    const int n0 = 0, nn = a.r.N;
    gn_bwd_reduce_kernel<T, KIND><<<dim3(slots, nn), NTHR, 0, s>>>(a, rows_pb, n0);
    MRISR_LAUNCH_CHECK();
    gn_bwd_finalize_kernel<T, KIND><<<nn * kGroups, FTHR, 0, s>>>(a, slots, inv_count, n0, nn, 0);
    MRISR_LAUNCH_CHECK();
    gn_bwd_apply_kernel<T, KIND><<<dim3(ceil_div(UH, rows_ap), nn), NTHR, 0, s>>>(a, rows_ap, n0);
    MRISR_LAUNCH_CHECK();
    gn_bwd_finalize_kernel<T, KIND><<<pcs, FTHR, 0, s>>>(a, slots, inv_count, 0, a.r.N, 1);
    MRISR_LAUNCH_CHECK();
    return 0;
}
}  // namespace

int gn_bwd_slots(int H, int W, int C) {
    // upper bound over both unit shapes (pool kinds use half the rows and half the columns)
    const int full = ceil_div(H, rows_per_block(W, C / 8, NTHR, UNITS_PER_THREAD));
    const int pool = ceil_div(H / 2, rows_per_block(W / 2, C / 8, NTHR, UNITS_PER_THREAD));
    const int m = full > pool ? full : pool;
    return m > 16 ? m : 16;                  // the fused kernel uses one slot per cluster rank (<= 16)
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
