// GroupNorm statistics / apply (+LeakyReLU, +2x2 max-pool), bilinear x2 up-sampling and its adjoint.
// Forward semantics follow models/unet_model.py:29-37 (conv -> GroupNorm(8) -> LeakyReLU(0.2)),
// :52 (MaxPool2d(2)), :71/:151 (Upsample(2, bilinear, align_corners=True)).
#include <stdlib.h>
#include <algorithm>
#include <map>
#include <mutex>
#include <vector>
#include "kernels.cuh"
#include "upsample_slide.cuh"

namespace mrisr {
namespace {
constexpr int NTHR = 256;
constexpr int PIX_PER_THREAD = 16;

template <typename T>
__global__ void pack_input_kernel(const float* __restrict__ x, Act<T> dst) {
    const int64_t total = (int64_t)dst.N * dst.H * dst.W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int xx = (int)(i % dst.W), yy = (int)((i / dst.W) % dst.H), n = (int)(i / ((int64_t)dst.W * dst.H));
        *dst.at(n, yy, xx, 0) = from_f<T>(x[i]);
    }
}

// ---- statistics: per (n, slot, group) partial sum / sum of squares -------------------------
// One block = `rows_pb` rows of one sample; thread mapping per common.cuh RowMap.
template <typename T>
__global__ void __launch_bounds__(NTHR) gn_stats_kernel(Act<T> r, float* __restrict__ part, int rows_pb) {
    __shared__ float s_red[2][NTHR * 8];
    __shared__ float s_ch[2][2048];
    const int C = r.C, nv = C / 8;
    const int tid = threadIdx.x;
    const int n = blockIdx.y;
    const RowMap m = row_map(r, NTHR, tid);
    float s1[8], s2[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) s1[k] = s2[k] = 0.f;
    if (m.active) {
        const int y_end = min(r.H, (int)(blockIdx.x + 1) * rows_pb);
        for (int y = blockIdx.x * rows_pb; y < y_end; ++y)
            for (int x = m.x0; x < r.W; x += m.xstep) {
                float v[8];
                load8(r.at(n, y, x, m.cv * 8), v);
#pragma unroll
                for (int k = 0; k < 8; ++k) { s1[k] += v[k]; s2[k] += v[k] * v[k]; }
            }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) { s_red[0][tid * 8 + k] = s1[k]; s_red[1][tid * 8 + k] = s2[k]; }
    __syncthreads();
    const bool planar = r.planar();
    const int cnt = NTHR / nv;
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
        part[(((int64_t)n * gridDim.x + blockIdx.x) * kGroups + g) * 2 + which] = t;
    }
}

__global__ void gn_finalize_kernel(const float* __restrict__ part, float* __restrict__ stats, int slots, double count) {
    const int ng = blockIdx.x;                   // n * 8 + g
    const int n = ng / kGroups, g = ng % kGroups;
    double s1 = 0.0, s2 = 0.0;
    for (int s = threadIdx.x; s < slots; s += 32) {
        const float* q = part + (((int64_t)n * slots + s) * kGroups + g) * 2;
        s1 += (double)q[0];
        s2 += (double)q[1];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
        s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    if (threadIdx.x == 0) {
        const double mean = s1 / count;
        double var = s2 / count - mean * mean;
        if (var < 0.0) var = 0.0;
        stats[ng * 2] = (float)mean;
        stats[ng * 2 + 1] = (float)(1.0 / sqrt(var + (double)kGnEps));
    }
}

// ---- apply: a = lrelu(gamma * (r - mean) * rstd + beta), optional 2x2 max-pool --------------
// One block = `rows_pb` (window) rows of one sample; a thread keeps one channel vector (scale/shift in registers).
template <typename T, bool POOL>
__global__ void __launch_bounds__(NTHR) gn_apply_kernel(Act<T> r, const float* __restrict__ stats,
                                                        const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, Act<T> a, Act<T> pooled, int rows_pb) {
    const int C = r.C, gs = C / kGroups;
    const int WH = POOL ? r.H / 2 : r.H, WW = POOL ? r.W / 2 : r.W;
    const int n = blockIdx.y;
    const RowMap m = row_map(r, NTHR, threadIdx.x);
    if (!m.active) return;
    const int cv = m.cv;
    float sc[8], sh[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int c = cv * 8 + k, g = c / gs;
        const float mean = stats[(n * kGroups + g) * 2], rstd = stats[(n * kGroups + g) * 2 + 1];
        sc[k] = gamma[c] * rstd;
        sh[k] = beta[c] - mean * sc[k];
    }
    const int wy_end = min(WH, (int)(blockIdx.x + 1) * rows_pb);
    // all loads of a batch are issued before the first store (the compiler cannot move a load above a store through
    // pointers that may alias): UB units (or one 2x2 window) in flight per thread
    constexpr int UB = 4;
    for (int wy = blockIdx.x * rows_pb; wy < wy_end; ++wy) {
        if (!POOL) {
            for (int wx0 = m.x0; wx0 < WW; wx0 += UB * m.xstep) {
                float v[UB][8];
#pragma unroll
                for (int u = 0; u < UB; ++u) load8(r.at(n, wy, min(wx0 + u * m.xstep, WW - 1), cv * 8), v[u]);
#pragma unroll
                for (int u = 0; u < UB; ++u) {
                    const int wx = wx0 + u * m.xstep;
                    if (wx >= WW) continue;
#pragma unroll
                    for (int k = 0; k < 8; ++k) v[u][k] = lrelu(fmaf(v[u][k], sc[k], sh[k]));
                    store8(a.at(n, wy, wx, cv * 8), v[u]);
                }
            }
        } else {
            for (int wx = m.x0; wx < WW; wx += m.xstep) {
                float v[4][8], mx[8];
#pragma unroll
                for (int q = 0; q < 4; ++q) load8(r.at(n, 2 * wy + (q >> 1), 2 * wx + (q & 1), cv * 8), v[q]);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) v[q][k] = lrelu(fmaf(v[q][k], sc[k], sh[k]));
                    store8(a.at(n, 2 * wy + (q >> 1), 2 * wx + (q & 1), cv * 8), v[q]);
                    // max over the values as stored (rounded to T) so backward's argmax recomputation agrees
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const float vr = to_f(from_f<T>(v[q][k]));
                        mx[k] = q == 0 ? vr : fmaxf(mx[k], vr);
                    }
                }
                store8(pooled.at(n, wy, wx, cv * 8), mx);
            }
        }
    }
}

// ---- bilinear x2, align_corners=True ------------------------------------------------------
// One block = `rows_pb` destination rows of one sample; mapped by the DESTINATION layout.
template <typename T>
__global__ void __launch_bounds__(NTHR) upsample2x_kernel(Act<T> src, Act<T> dst, float sy, float sx, int rows_pb) {
    const int n = blockIdx.y;
    const RowMap m = row_map(dst, NTHR, threadIdx.x);
    if (!m.active) return;
    const int cv = m.cv;
    const int y_end = min(dst.H, (int)(blockIdx.x + 1) * rows_pb);
    for (int Y = blockIdx.x * rows_pb; Y < y_end; ++Y) {
        const Lerp ly = up_coord(Y, src.H, sy);
        const float h1 = ly.w1, h0 = 1.f - h1;
        constexpr int UB = 2;                          // destination pixels in flight per thread (8 loads before the first store)
        for (int X0 = m.x0; X0 < dst.W; X0 += UB * m.xstep) {
            float v[UB][4][8], w1[UB];
#pragma unroll
            for (int u = 0; u < UB; ++u) {
                const Lerp lx = up_coord(min(X0 + u * m.xstep, dst.W - 1), src.W, sx);
                w1[u] = lx.w1;
                load8(src.at(n, ly.i0, lx.i0, cv * 8), v[u][0]);
                load8(src.at(n, ly.i0, lx.i1, cv * 8), v[u][1]);
                load8(src.at(n, ly.i1, lx.i0, cv * 8), v[u][2]);
                load8(src.at(n, ly.i1, lx.i1, cv * 8), v[u][3]);
            }
#pragma unroll
            for (int u = 0; u < UB; ++u) {
                const int X = X0 + u * m.xstep;
                if (X >= dst.W) continue;
                const float a1 = w1[u], a0 = 1.f - a1;
                float o[8];
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    o[k] = h0 * (a0 * v[u][0][k] + a1 * v[u][1][k]) + h1 * (a0 * v[u][2][k] + a1 * v[u][3][k]);
                store8(dst.at(n, Y, X, cv * 8), o);
            }
        }
    }
}

__device__ __forceinline__ void adj_range(int i, int in_size, float scale, int& lo, int& hi) {
    const int out_size = 2 * in_size;
    if (scale <= 0.f) { lo = 0; hi = out_size - 1; return; }
    lo = (int)floorf((float)(i - 1) / scale) - 1;
    hi = (int)ceilf((float)(i + 1) / scale) + 1;
    if (lo < 0) lo = 0;
    if (hi > out_size - 1) hi = out_size - 1;
}
__device__ __forceinline__ float adj_weight(int dst, int i, int in_size, float scale) {
    const Lerp l = up_coord(dst, in_size, scale);
    return (l.i0 == i ? 1.f - l.w1 : 0.f) + (l.i1 == i ? l.w1 : 0.f);
}

// Candidate destination indices of source index i and their weights: dst d contributes to i with (1 - w1) if i0(d) == i and
// with w1 if i1(d) == i.  With scale = (in-1)/(2in-1) < 1/2 at most 5 consecutive d qualify.
struct AdjTaps { int lo, cnt; float w[6]; };
__device__ __forceinline__ AdjTaps adj_taps(int i, int in_size, float scale) {
    AdjTaps t;
    int lo, hi;
    adj_range(i, in_size, scale, lo, hi);
    // shrink to the first / last non-zero weight (the range above is a safe superset)
    while (lo <= hi && adj_weight(lo, i, in_size, scale) == 0.f) ++lo;
    while (hi >= lo && adj_weight(hi, i, in_size, scale) == 0.f) --hi;
    t.lo = lo;
    t.cnt = hi - lo + 1;
    if (t.cnt > 6) t.cnt = 6;                       // cannot happen for scale < 1/2 (kept as a bound for the register array)
#pragma unroll
    for (int j = 0; j < 6; ++j) t.w[j] = j < t.cnt ? adj_weight(lo + j, i, in_size, scale) : 0.f;
    return t;
}

// gather form of the adjoint: deterministic, no atomics.  One block = `rows_pb` source rows of one sample, mapped by the
// layout of the (4x larger) gradient tensor being read.  The tap tables (first destination index, count, weights) of every
// source column and of the block's rows are built once per block in shared memory: computing them per pixel cost more
// instructions than the gather itself.
constexpr int kAdjMaxW = 1024, kAdjMaxRows = 64;
struct AdjTap5 { int lo, cnt; float w[5]; float pad; };           // 32 B
template <typename T>
__global__ void __launch_bounds__(NTHR) upsample2x_bwd_kernel(Act<T> ddst, Act<T> addend, Act<T> dsrc, float sy,
                                                              float sx, int rows_pb) {
    __shared__ AdjTap5 s_tx[kAdjMaxW];
    __shared__ AdjTap5 s_ty[kAdjMaxRows];
    const int n = blockIdx.y;
    const int y_begin = blockIdx.x * rows_pb, y_end = min(dsrc.H, y_begin + rows_pb);
    for (int i = threadIdx.x; i < dsrc.W + (y_end - y_begin); i += NTHR) {
        const bool is_x = i < dsrc.W;
        const AdjTaps t = is_x ? adj_taps(i, dsrc.W, sx) : adj_taps(y_begin + i - dsrc.W, dsrc.H, sy);
        AdjTap5& o = is_x ? s_tx[i] : s_ty[i - dsrc.W];
        o.lo = t.lo;
        o.cnt = t.cnt < 5 ? t.cnt : 5;
#pragma unroll
        for (int j = 0; j < 5; ++j) o.w[j] = t.w[j];
    }
    __syncthreads();
    const RowMap m = row_map(ddst, NTHR, threadIdx.x);
    if (!m.active) return;
    const int cv = m.cv;
    const int ps = ddst.ps ? ddst.ps : ddst.C;
    for (int y = y_begin; y < y_end; ++y) {
        const AdjTap5 ty = s_ty[y - y_begin];
        for (int x = m.x0; x < dsrc.W; x += m.xstep) {
            const AdjTap5 tx = s_tx[x];
            float acc[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] = 0.f;
            if (addend.p) load8(addend.at(n, y, x, cv * 8), acc);
#pragma unroll
            for (int jy = 0; jy < 5; ++jy) {
                if (jy >= ty.cnt) continue;
                const T* rowp = ddst.at(n, ty.lo + jy, tx.lo, cv * 8);
                float h[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) h[k] = 0.f;
#pragma unroll
                for (int jx = 0; jx < 5; ++jx) {
                    if (jx < tx.cnt) {
                        float v[8];
                        load8(rowp + (int64_t)jx * ps, v);
#pragma unroll
                        for (int k = 0; k < 8; ++k) h[k] = fmaf(tx.w[jx], v[k], h[k]);
                    }
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) acc[k] = fmaf(ty.w[jy], h[k], acc[k]);
            }
            store8(dsrc.at(n, y, x, cv * 8), acc);
        }
    }
}

// Tiled form for bf16: a CTA owns TR source rows x TW source columns of one (sample, channel-vector range).  The destination
// gradient rows / columns that tile gathers from are contiguous 16-byte vectors in both layouts (one plane of a planar-8
// tensor, or all channels of an NHWC tensor), so they are staged global -> shared memory once with cp.async (coalesced,
// everything in flight at once) and the ~16-20 taps of every output then read shared memory instead of L1/L2.
// (A row-ring form -- a CTA walking down source rows with the destination rows streaming through a 12-slot cp.async ring, every
// row fetched once -- was measured in round 2 and lost, 0.634 -> 0.825 ms at 32@256^2: two barriers per source row with eight
// warps per CTA cost more than the tile's load-wait-compute bubble; profiles/r02_lost_variants.md.)
__device__ __forceinline__ void cp_async16_na(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
struct UpBwdTile { int TR, TW, ncv, rows_cap, cols_cap, tiles_x, tiles_y; };

// tap table of one axis (in_size entries), built once per size and kept for the life of the library
__global__ void adj_table_kernel(AdjTap5* __restrict__ tab, int in_size, float scale) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= in_size) return;
    const AdjTaps t = adj_taps(i, in_size, scale);
    AdjTap5 o;
    o.lo = t.lo;
    o.cnt = t.cnt < 5 ? t.cnt : 5;
#pragma unroll
    for (int j = 0; j < 5; ++j) o.w[j] = t.w[j];
    o.pad = 0.f;
    tab[i] = o;
}

__global__ void __launch_bounds__(NTHR, 2) upsample2x_bwd_tiled_kernel(Act<__nv_bfloat16> ddst, Act<__nv_bfloat16> addend,
                                                                       Act<__nv_bfloat16> dsrc, const AdjTap5* __restrict__ tab_y,
                                                                       const AdjTap5* __restrict__ tab_x, UpBwdTile tp) {
    extern __shared__ uint4 up_smem[];
    uint4* rows = up_smem;                                                          // [rows_cap][cols_cap] 16-byte vectors
    AdjTap5* s_ty = reinterpret_cast<AdjTap5*>(up_smem + (size_t)tp.rows_cap * tp.cols_cap);    // [TR]
    const int tid = threadIdx.x, ncv = tp.ncv;
    const int tile_x = blockIdx.x % tp.tiles_x, tile_y = blockIdx.x / tp.tiles_x;
    const int units_per_sample = (dsrc.C / 8) / ncv;
    const int n = blockIdx.y / units_per_sample, cv0 = (blockIdx.y % units_per_sample) * ncv;
    const int y0 = tile_y * tp.TR, x0 = tile_x * tp.TW;
    const int th = min(tp.TR, dsrc.H - y0), tw = min(tp.TW, dsrc.W - x0);
    const AdjTap5 ya = tab_y[y0], yb = tab_y[y0 + th - 1], xa = tab_x[x0], xb = tab_x[x0 + tw - 1];
    const int d_lo = ya.lo, nrows = yb.lo + yb.cnt - d_lo;
    const int X_lo = xa.lo, ncols = (xb.lo + xb.cnt - X_lo) * ncv;
    if (nrows > tp.rows_cap || ncols > tp.cols_cap) __trap();                       // the host sized the caps from the same taps
    for (int r = 0; r < nrows; ++r) {
        const __nv_bfloat16* src = ddst.at(n, d_lo + r, X_lo, cv0 * 8);
        uint4* dst = rows + (size_t)r * tp.cols_cap;
        for (int c = tid; c < ncols; c += NTHR) cp_async16_na(dst + c, src + (int64_t)c * 8);
    }
    if (tid < th * 2) cp_async16_na(reinterpret_cast<uint4*>(s_ty) + tid, reinterpret_cast<const uint4*>(tab_y + y0) + tid);
    asm volatile("cp.async.commit_group;" ::: "memory");
    // a thread owns at most two (column, channel vector) pairs of the tile and walks the tile's rows: its horizontal taps stay
    // in registers
    const int per_row = tw * ncv;
    AdjTap5 tx[2];
    int xl[2], cvl[2], off[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const int j = min(tid + u * NTHR, per_row - 1);
        xl[u] = j / ncv;
        cvl[u] = j - xl[u] * ncv;
        tx[u] = tab_x[x0 + xl[u]];
        off[u] = (tx[u].lo - X_lo) * ncv + cvl[u];
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    for (int yl = 0; yl < th; ++yl) {
        const AdjTap5 ty = s_ty[yl];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            if (tid + u * NTHR >= per_row) continue;
            float add[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) add[k] = 0.f;
            if (addend.p) load8(addend.at(n, y0 + yl, x0 + xl[u], (cv0 + cvl[u]) * 8), add);   // consumed last
            const uint4* base = rows + (size_t)(ty.lo - d_lo) * tp.cols_cap + off[u];
            float acc[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] = 0.f;
            for (int jy = 0; jy < ty.cnt; ++jy) {                     // warp-uniform trip count (one row per iteration of yl)
                const float wy = s_ty[yl].w[jy];
                float h[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) h[k] = 0.f;
#pragma unroll
                for (int jx = 0; jx < 5; ++jx) {
                    if (jx < tx[u].cnt) {
                        const uint4 raw = base[(size_t)jy * tp.cols_cap + jx * ncv];
                        const uint32_t w4[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            h[2 * i] = fmaf(tx[u].w[jx], __uint_as_float(w4[i] << 16), h[2 * i]);
                            h[2 * i + 1] = fmaf(tx[u].w[jx], __uint_as_float(w4[i] & 0xffff0000u), h[2 * i + 1]);
                        }
                    }
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) acc[k] = fmaf(wy, h[k], acc[k]);
            }
            float o[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = add[k] + acc[k];
            store8(dsrc.at(n, y0 + yl, x0 + xl[u], (cv0 + cvl[u]) * 8), o);
        }
    }
}

// device tap tables, one per axis length, built on first use (a handful of sizes per network)
static const AdjTap5* adj_table(int in_size, cudaStream_t s) {
    static std::mutex mu;
    static std::map<int, AdjTap5*> tabs;
    std::lock_guard<std::mutex> lock(mu);
    auto it = tabs.find(in_size);
    if (it != tabs.end()) return it->second;
    AdjTap5* p = nullptr;
    if (cudaMalloc(&p, (size_t)in_size * sizeof(AdjTap5)) != cudaSuccess) return nullptr;
    adj_table_kernel<<<ceil_div(in_size, 128), 128, 0, s>>>(p, in_size, up_scale(in_size));
    count_launch();
    if (cudaStreamSynchronize(s) != cudaSuccess) { cudaFree(p); return nullptr; }   // first use only: other streams may read it next
    tabs[in_size] = p;
    return p;
}

// host replica of the tap ranges (same float arithmetic as up_coord / adj_weight): first and last destination index with a
// non-zero weight for every source index
static void adj_ranges_host(int in_size, std::vector<int>& lo, std::vector<int>& hi) {
    const float scale = up_scale(in_size);
    lo.assign(in_size, 1 << 30);
    hi.assign(in_size, -1);
    for (int d = 0; d < 2 * in_size; ++d) {
        const float sf = scale * (float)d;
        const int i0 = (int)sf, i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
        const float w1 = sf - (float)i0;
        const float wa = (1.f - w1) + (i1 == i0 ? w1 : 0.f);               // adj_weight(d, i0)
        if (wa != 0.f) { lo[i0] = std::min(lo[i0], d); hi[i0] = std::max(hi[i0], d); }
        if (i1 != i0 && w1 != 0.f) { lo[i1] = std::min(lo[i1], d); hi[i1] = std::max(hi[i1], d); }
    }
}
static int max_span(const std::vector<int>& lo, const std::vector<int>& hi, int tile) {
    int m = 0;
    const int n = (int)lo.size();
    for (int a = 0; a < n; a += tile) {
        const int b = std::min(n, a + tile) - 1;
        m = std::max(m, hi[b] - lo[a] + 1);
    }
    return m;
}

static bool up_bwd_plan(const Act<__nv_bfloat16>& ddst, const Act<__nv_bfloat16>& dsrc, UpBwdTile& tp, size_t& smem) {
    if (dsrc.C % 8) return false;
    if (ddst.planar() != dsrc.planar()) return false;      // level 1 -> level 0 boundary: per-vector strided stores, the gather form is faster
    const int nv = dsrc.C / 8;
    // the staged tensor is the (4x larger) destination gradient: units follow ITS layout (a plane, or all channels of an NHWC
    // tensor); the low-resolution result is addressed per vector and may have the other layout (level 1 -> level 0 boundary)
    tp.ncv = ddst.planar() ? 1 : nv;
    std::vector<int> ylo, yhi, xlo, xhi;
    adj_ranges_host(dsrc.H, ylo, yhi);
    adj_ranges_host(dsrc.W, xlo, xhi);
    for (int i = 0; i < dsrc.H; ++i) if (yhi[i] < 0 || yhi[i] - ylo[i] + 1 > 5) return false;
    for (int i = 0; i < dsrc.W; ++i) if (xhi[i] < 0 || xhi[i] - xlo[i] + 1 > 5) return false;
    int tw = (528 / tp.ncv - 4) / 2;                                  // tile columns x channel vectors <= 2 x 256 threads
    if (tw * tp.ncv > 2 * NTHR) tw = 2 * NTHR / tp.ncv;
    if (tw < 4) return false;
    if (tw > dsrc.W) tw = dsrc.W;
    tp.TW = tw;
    tp.cols_cap = max_span(xlo, xhi, tw) * tp.ncv;
    const size_t budget = 100u * 1024u;                               // two CTAs per SM
    tp.TR = 0;
    for (int tr = std::min(dsrc.H, 16); tr >= 1; --tr) {
        const int rc = max_span(ylo, yhi, tr);
        const size_t need = (size_t)rc * tp.cols_cap * 16 + (size_t)tr * sizeof(AdjTap5);
        if (need <= budget) { tp.TR = tr; tp.rows_cap = rc; smem = need; break; }
    }
    if (!tp.TR) return false;
    tp.tiles_x = ceil_div(dsrc.W, tp.TW);
    tp.tiles_y = ceil_div(dsrc.H, tp.TR);
    return true;
}

inline int grid_for(int64_t total) {
    int64_t b = ceil_div64(total, NTHR);
    if (b > 148 * 16) b = 148 * 16;
    if (b < 1) b = 1;
    return (int)b;
}
}  // namespace

template <typename T>
int launch_pack_input(const float* x, Act<T> dst, cudaStream_t s) {
    ProfScope prof_("pack_input", s);
    pack_input_kernel<T><<<grid_for((int64_t)dst.N * dst.H * dst.W), NTHR, 0, s>>>(x, dst);
    MRISR_LAUNCH_CHECK();
    return 0;
}

int stat_slots_generic(int H, int W, int C) {
    return ceil_div(H, rows_per_block(W, C / 8, NTHR, PIX_PER_THREAD));
}

template <typename T>
int launch_gn_stats(Act<T> r, float* part, cudaStream_t s) {
    ProfScope prof_("gn_stats", s);
    dim3 grid(stat_slots_generic(r.H, r.W, r.C), r.N);
    gn_stats_kernel<T><<<grid, NTHR, 0, s>>>(r, part, rows_per_block(r.W, r.C / 8, NTHR, PIX_PER_THREAD));
    MRISR_LAUNCH_CHECK();
    return 0;
}

int launch_gn_finalize(const float* part, float* stats, int N, int slots, double count, cudaStream_t s) {
    ProfScope prof_("gn_finalize", s);
    gn_finalize_kernel<<<N * kGroups, 32, 0, s>>>(part, stats, slots, count);
    MRISR_LAUNCH_CHECK();
    return 0;
}

template <typename T>
int launch_gn_apply(Act<T> r, const float* stats, const float* gamma, const float* beta, Act<T> a, Act<T> pooled,
                    cudaStream_t s) {
    const char* tag = "gn_apply";
    char tbuf[96];
    if (prof_on()) {
        snprintf(tbuf, sizeof(tbuf), "gn_apply:pool%d:%d@%dx%d", pooled.p ? 1 : 0, r.C, r.H, r.W);
        tag = prof_intern(tbuf);
    }
    ProfScope prof_(tag, s);
    const int nv = r.C / 8;
    // vectors per thread: 16 (plain) / 4 (with the pooled quarter).  Swept on B200 in round 2 (32, 64 / 8): no gain (0.99 ->
    // 1.01, 1.08 / 0.99 ms per step), the kernels run at ~83 % of the HBM copy rate.
    constexpr int upt = 16, upt_pool = 4;
    if (pooled.p) {
        const int rpb = rows_per_block(r.W / 2, nv, NTHR, upt_pool);
        dim3 grid(ceil_div(r.H / 2, rpb), r.N);
        gn_apply_kernel<T, true><<<grid, NTHR, 0, s>>>(r, stats, gamma, beta, a, pooled, rpb);
    } else {
        const int rpb = rows_per_block(r.W, nv, NTHR, upt);
        dim3 grid(ceil_div(r.H, rpb), r.N);
        gn_apply_kernel<T, false><<<grid, NTHR, 0, s>>>(r, stats, gamma, beta, a, pooled, rpb);
    }
    MRISR_LAUNCH_CHECK();
    return 0;
}

// forward bilinear x2 that also leaves the GroupNorm partials of dst in `part` (slots = the returned *slots per sample);
// returns MRISR_NOT_APPLICABLE when the shape needs the plain path (the caller then runs gn_stats itself)
template <typename T>
int launch_upsample2x_stats(Act<T> src, Act<T> dst, float* part, size_t part_cap_floats, int* slots, cudaStream_t s) {
    if (!((int64_t)(src.W + 2) * (src.ps ? src.ps : src.C) < (1ll << 31) / (src.H + 2) &&
          (int64_t)(dst.W + 2) * (dst.ps ? dst.ps : dst.C) < (1ll << 31) / (dst.H + 2)) || dst.C > 2048 || dst.C % 8)
        return MRISR_NOT_APPLICABLE;
    const int rows = dst.H >= 64 ? 16 : 8;
    dim3 grid_s(ceil_div(dst.H, rows), dst.N);
    if ((size_t)dst.N * grid_s.x * kGroups * 2 > part_cap_floats) return MRISR_NOT_APPLICABLE;
    ProfScope prof_("upsample2x", s);
    upsample2x_slide_stats_kernel<T, NTHR><<<grid_s, NTHR, 0, s>>>(src, dst, up_scale(src.H), up_scale(src.W), rows, part);
    MRISR_LAUNCH_CHECK();
    *slots = (int)grid_s.x;
    return 0;
}

template <typename T>
int launch_upsample2x(Act<T> src, Act<T> dst, cudaStream_t s) {
    ProfScope prof_("upsample2x", s);
    // columns stay with a thread, source rows slide through registers (upsample_slide.cuh); 16 destination rows per block
    // re-read 1 - 2 of their 9 - 10 source rows from the neighbouring blocks' (L2).  Measured on B200 (round 2, batch 64
    // train step): 0.706 -> 0.556 ms per step against the gather form below, which stays for tensors beyond 32-bit offsets.
    if ((int64_t)(src.W + 2) * (src.ps ? src.ps : src.C) < (1ll << 31) / (src.H + 2) &&
        (int64_t)(dst.W + 2) * (dst.ps ? dst.ps : dst.C) < (1ll << 31) / (dst.H + 2)) {
        const int rows = dst.H >= 64 ? 16 : 8;
        dim3 grid_s(ceil_div(dst.H, rows), dst.N);
        upsample2x_slide_kernel<T, 1, NTHR><<<grid_s, NTHR, 0, s>>>(src, dst, up_scale(src.H), up_scale(src.W), rows);
        MRISR_LAUNCH_CHECK();
        return 0;
    }
    const int rpb = rows_per_block(dst.W, dst.C / 8, NTHR, 8);
    dim3 grid(ceil_div(dst.H, rpb), dst.N);
    upsample2x_kernel<T><<<grid, NTHR, 0, s>>>(src, dst, up_scale(src.H), up_scale(src.W), rpb);
    MRISR_LAUNCH_CHECK();
    return 0;
}

template <typename T>
int launch_upsample2x_bwd(Act<T> ddst, Act<T> addend, Act<T> dsrc, cudaStream_t s) {
    const char* tag = "upsample2x_bwd";
    char tbuf[96];
    if (prof_on()) {
        snprintf(tbuf, sizeof(tbuf), "upsample2x_bwd:add%d:%d@%dx%d", addend.p ? 1 : 0, dsrc.C, dsrc.H, dsrc.W);
        tag = prof_intern(tbuf);
    }
    ProfScope prof_(tag, s);
    if constexpr (sizeof(T) == 2) {
        UpBwdTile tp;
        size_t smem = 0;
        if (up_bwd_plan(ddst, dsrc, tp, smem)) {
            static PerDeviceFlag attr_done_pd; bool& attr_done = attr_done_pd.cur();
            if (!attr_done) {
                MRISR_CUDA_CHECK(cudaFuncSetAttribute(upsample2x_bwd_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
                attr_done = true;
            }
            const AdjTap5* ty = adj_table(dsrc.H, s);
            const AdjTap5* tx = dsrc.W == dsrc.H ? ty : adj_table(dsrc.W, s);
            if (!ty || !tx) { set_error("upsample2x_bwd: tap table allocation failed"); return -1; }
            dim3 grid(tp.tiles_x * tp.tiles_y, dsrc.N * ((dsrc.C / 8) / tp.ncv));
            upsample2x_bwd_tiled_kernel<<<grid, NTHR, smem, s>>>(ddst, addend, dsrc, ty, tx, tp);
            MRISR_LAUNCH_CHECK();
            return 0;
        }
    }
    int rpb = rows_per_block(dsrc.W, dsrc.C / 8, NTHR, 4);
    if (rpb > kAdjMaxRows) rpb = kAdjMaxRows;
    if (dsrc.W > kAdjMaxW) { set_error("upsample2x_bwd: W = %d exceeds the tap table (%d)", dsrc.W, kAdjMaxW); return -1; }
    dim3 grid(ceil_div(dsrc.H, rpb), dsrc.N);
    upsample2x_bwd_kernel<T><<<grid, NTHR, 0, s>>>(ddst, addend, dsrc, up_scale(dsrc.H), up_scale(dsrc.W), rpb);
    MRISR_LAUNCH_CHECK();
    return 0;
}

#define INST(T)                                                                                                   \
    template int launch_pack_input<T>(const float*, Act<T>, cudaStream_t);                                        \
    template int launch_gn_stats<T>(Act<T>, float*, cudaStream_t);                                                \
    template int launch_gn_apply<T>(Act<T>, const float*, const float*, const float*, Act<T>, Act<T>, cudaStream_t); \
    template int launch_upsample2x<T>(Act<T>, Act<T>, cudaStream_t);                                              \
    template int launch_upsample2x_stats<T>(Act<T>, Act<T>, float*, size_t, int*, cudaStream_t);                  \
    template int launch_upsample2x_bwd<T>(Act<T>, Act<T>, Act<T>, cudaStream_t);
INST(float)
INST(__nv_bfloat16)
#undef INST

}  // namespace mrisr
