// Bilinear x2 (align_corners=True) with the two source rows of a destination column kept in registers (the default forward
// kernel since round 2; upsample2x_kernel in norm_act.cu is the fallback for tensors beyond 32-bit offsets).
// STATS: the thread also sums the values it stores (as rounded to T) and their squares per channel; the kernel reduces them to
// per-(sample, block, group) GroupNorm partials, so the up-branch tensors need no separate gn_stats pass (round 2).
//
// upsample2x_kernel gathers 4 source vectors and evaluates 5 Act::at address chains per destination vector: 325 SASS
// instructions per two destination vectors, issue-bound at ~42 % of the HBM copy rate (profiles/r01_elementwise_issue.md).
// Here a thread owns destination columns (fixed horizontal taps and weights, base pointers computed once) and walks the
// block's destination rows: consecutive rows share at least one source row, so each step needs at most one new source row
// (2 loads, horizontal blend) and one vertical blend.  Same expression tree as the default kernel:
//   out = h0 * (a0 * s[i0y][i0x] + a1 * s[i0y][i1x]) + h1 * (a0 * s[i1y][i0x] + a1 * s[i1y][i1x]).
// The body is a plain function of (sample, block, thread) so that tests/native/upsample_slide_host.cu runs it on the CPU.
#pragma once
#include "common.cuh"

namespace mrisr {

// 8-channel vector held as loaded (4 registers for bf16): what the prefetch keeps in flight; unpack_raw gives load8's values
template <typename T> struct RawVec;
template <> struct RawVec<float> { float4 a, b; };
template <> struct RawVec<__nv_bfloat16> { uint4 a; };
__host__ __device__ __forceinline__ void load_raw(const float* p, RawVec<float>& r) {
    r.a = *reinterpret_cast<const float4*>(p);
    r.b = *reinterpret_cast<const float4*>(p + 4);
}
__host__ __device__ __forceinline__ void load_raw(const __nv_bfloat16* p, RawVec<__nv_bfloat16>& r) { r.a = *reinterpret_cast<const uint4*>(p); }
__host__ __device__ __forceinline__ void unpack_raw(const RawVec<float>& r, float (&v)[8]) {
    v[0] = r.a.x; v[1] = r.a.y; v[2] = r.a.z; v[3] = r.a.w; v[4] = r.b.x; v[5] = r.b.y; v[6] = r.b.z; v[7] = r.b.w;
}
__host__ __device__ __forceinline__ void unpack_raw(const RawVec<__nv_bfloat16>& r, float (&v)[8]) {
    v[0] = bits_to_float(r.a.x << 16); v[1] = bits_to_float(r.a.x & 0xffff0000u);
    v[2] = bits_to_float(r.a.y << 16); v[3] = bits_to_float(r.a.y & 0xffff0000u);
    v[4] = bits_to_float(r.a.z << 16); v[5] = bits_to_float(r.a.z & 0xffff0000u);
    v[6] = bits_to_float(r.a.w << 16); v[7] = bits_to_float(r.a.w & 0xffff0000u);
}

// PF: the source row the NEXT destination row will need is loaded (raw) before the current row is blended and stored, so its
// latency overlaps that work instead of stalling the next iteration (16 more registers per column).
template <typename T, int UB, bool PF = false, bool STATS = false>
__host__ __device__ __forceinline__ void upsample2x_slide_body(const Act<T>& src, const Act<T>& dst, float sy, float sx, int rows_pb,
                                                               int n, int block, int tid, int nthr, float* s1 = nullptr,
                                                               float* s2 = nullptr) {
    const RowMap m = row_map(dst, nthr, tid);
    if (!m.active) return;
    const int Y0 = block * rows_pb, Y1 = Y0 + rows_pb < dst.H ? Y0 + rows_pb : dst.H;
    const int pss = src.ps ? src.ps : src.C, psd = dst.ps ? dst.ps : dst.C;
    const int srs = (src.W + 2) * pss, drs = (dst.W + 2) * psd;           // row strides in elements (< 2^31: host check)
    for (int X0 = m.x0; X0 < dst.W; X0 += UB * m.xstep) {
        const T* sp[UB];
        T* dp[UB];
        int d1[UB];
        float a1[UB];
        bool on[UB];
MRISR_UNROLL
        for (int u = 0; u < UB; ++u) {
            const int X = X0 + u * m.xstep;
            on[u] = X < dst.W;
            const int Xc = on[u] ? X : dst.W - 1;                         // clamped duplicate: loaded, never stored
            const Lerp lx = up_coord(Xc, src.W, sx);
            a1[u] = lx.w1;
            sp[u] = src.at(n, 0, lx.i0, m.cv * 8);
            d1[u] = (lx.i1 - lx.i0) * pss;
            dp[u] = dst.at(n, 0, Xc, m.cv * 8);
        }
        float top[UB][8], bot[UB][8];
        int ti = -1, bi = -1;                                             // source rows held in top / bot
        RawVec<T> p0[PF ? UB : 1], p1[PF ? UB : 1];                       // PF: raw taps of source row pr, loaded one step ahead
        int pr = -1;
        auto fetch_row = [&](int r) {
MRISR_UNROLL
            for (int u = 0; u < (PF ? UB : 1); ++u) {
                load_raw(sp[u] + (int64_t)r * srs, p0[u]);
                load_raw(sp[u] + (int64_t)r * srs + d1[u], p1[u]);
            }
        };
        auto load_row = [&](int r, float (&out)[UB][8]) {
            float v0[UB][8], v1[UB][8];
            if (PF && r == pr) {
MRISR_UNROLL
                for (int u = 0; u < UB; ++u) { unpack_raw(p0[PF ? u : 0], v0[u]); unpack_raw(p1[PF ? u : 0], v1[u]); }
            } else {
MRISR_UNROLL
                for (int u = 0; u < UB; ++u) {
                    load8(sp[u] + (int64_t)r * srs, v0[u]);
                    load8(sp[u] + (int64_t)r * srs + d1[u], v1[u]);
                }
            }
MRISR_UNROLL
            for (int u = 0; u < UB; ++u) {
                const float a0 = 1.f - a1[u];
MRISR_UNROLL
                for (int k = 0; k < 8; ++k) out[u][k] = a0 * v0[u][k] + a1[u] * v1[u][k];
            }
        };
        auto copy_row = [&](const float (&from)[UB][8], float (&to)[UB][8]) {
MRISR_UNROLL
            for (int u = 0; u < UB; ++u)
MRISR_UNROLL
                for (int k = 0; k < 8; ++k) to[u][k] = from[u][k];
        };
        for (int Y = Y0; Y < Y1; ++Y) {
            const Lerp ly = up_coord(Y, src.H, sy);
            if (ly.i0 != ti) {
                if (ly.i0 == bi) copy_row(bot, top); else load_row(ly.i0, top);
                ti = ly.i0;
            }
            if (ly.i1 != bi) {
                if (ly.i1 == ti) copy_row(top, bot); else load_row(ly.i1, bot);
                bi = ly.i1;
            }
            if (PF && Y + 1 < Y1) {                                       // the row the next step will have to load, if any
                const Lerp ln = up_coord(Y + 1, src.H, sy);
                const int want = ln.i0 != ti && ln.i0 != bi ? ln.i0 : (ln.i1 != bi && ln.i1 != ti && ln.i1 != ln.i0 ? ln.i1 : -1);
                if (want >= 0 && want != pr) { fetch_row(want); pr = want; }
            }
            const float h1 = ly.w1, h0 = 1.f - h1;
MRISR_UNROLL
            for (int u = 0; u < UB; ++u) {
                if (!on[u]) continue;
                float o[8];
MRISR_UNROLL
                for (int k = 0; k < 8; ++k) o[k] = h0 * top[u][k] + h1 * bot[u][k];
                store8(dp[u] + (int64_t)Y * drs, o);
                if (STATS) {
MRISR_UNROLL
                    for (int k = 0; k < 8; ++k) {
                        const float q = to_f(from_f<T>(o[k]));             // statistics of the tensor as stored
                        s1[k] += q;
                        s2[k] += q * q;
                    }
                }
            }
        }
    }
}

#ifdef __CUDACC__
template <typename T, int UB, int NT, bool PF = false>
__global__ void __launch_bounds__(NT) upsample2x_slide_kernel(Act<T> src, Act<T> dst, float sy, float sx, int rows_pb) {
    upsample2x_slide_body<T, UB, PF>(src, dst, sy, sx, rows_pb, blockIdx.y, blockIdx.x, threadIdx.x, NT);
}

// + GroupNorm partial statistics of dst: part[((n * gridDim.x + block) * 8 + g) * 2 + {sum, sum of squares}] (gn_stats_kernel's
// layout with slots = gridDim.x, finalised by gn_finalize_kernel)
template <typename T, int NT>
__global__ void __launch_bounds__(NT) upsample2x_slide_stats_kernel(Act<T> src, Act<T> dst, float sy, float sx, int rows_pb,
                                                                    float* __restrict__ part) {
    __shared__ float s_red[2][NT * 8];
    __shared__ float s_ch[2][2048];
    float s1[8], s2[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) s1[k] = s2[k] = 0.f;
    upsample2x_slide_body<T, 1, false, true>(src, dst, sy, sx, rows_pb, blockIdx.y, blockIdx.x, threadIdx.x, NT, s1, s2);
    const int tid = threadIdx.x, C = dst.C, nv = C / 8;
#pragma unroll
    for (int k = 0; k < 8; ++k) { s_red[0][tid * 8 + k] = s1[k]; s_red[1][tid * 8 + k] = s2[k]; }
    __syncthreads();
    const bool planar = dst.planar();
    const int cnt = NT / nv;                      // inactive threads (NT % nv != 0) hold zeros
    for (int i = tid; i < 2 * C; i += NT) {
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
        part[(((int64_t)blockIdx.y * gridDim.x + blockIdx.x) * kGroups + g) * 2 + which] = t;
    }
}
#endif

}  // namespace mrisr
