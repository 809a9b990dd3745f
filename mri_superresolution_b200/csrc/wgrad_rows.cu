// tcgen05 weight gradient for the narrow high-resolution layers, row-structured (second generation of wgrad_ring.cu):
//   dW[tap][ci][co] = sum_p X[p + shift(tap)][ci] * dY[p][co]           (autograd of nn.Conv2d, models/unet_model.py:29..168)
// wgrad_ring.cu issues one M=64 MMA per (tap, 16 pixels): 72 tiny MMAs per 128 pixels, ~50 cycles each - 10x over the HBM
// time of these layers.  Here the three VERTICAL taps are folded into the M dimension of one MMA at no data-movement cost:
// image-row segments of X live in shared memory as  [row slot][8-channel plane][pixels x 16 B]  with
// row-slot stride = planes * plane stride, so the M groups (dy, plane) of three consecutive rows form ONE arithmetic
// progression of shared-memory addresses - exactly what the MN-major operand descriptor expresses (SBO = plane stride).
// The horizontal taps stay separate MMAs (descriptor start + dx * 16 B).  Per 128 pixels: 24 MMAs of M = 64 / 128 instead
// of 72 of M = 64, every X byte still fetched once (+2 halo rows per row chunk), dY once.
//   A (M side) = X rows y-1, y, y+1: M = 3 * planes * 8 channels (48 -> 64, 96 -> 128; padding groups read the next slot)
//   B (N side) = dY row y: N = Cout;  K = pixels, 16 per MMA;  D = 3 accumulators (dx) x [M x Cout] fp32 in TMEM.
// A CTA walks work units (sample, column strip <= 256 wide, chunk of rows), keeps accumulating into the same TMEM tile and
// writes one partial at the end; wgrad_rows_reduce_kernel sums the partials in a fixed order (deterministic).
// Warp roles: warp 0 producer (cp.async.bulk per plane row segment), warp 1 MMA issuer, warps 2..5 epilogue.
#include <stdlib.h>
#include "kernels.cuh"
#include "tc_common.cuh"

namespace mrisr {
namespace tc {

int num_sms();

__device__ __forceinline__ void bulk_load_r(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ bool elect_one_r() {
    uint32_t pred;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "elect.sync _|p, 0xffffffff;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void umma_lohi_r(uint32_t tmem_d, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi,
                                            uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b64 da, db;\n"
        "setp.ne.b32 p, %6, 0;\n"
        "mov.b64 da, {%1, %2};\n"
        "mov.b64 db, {%3, %4};\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n"
        "}\n"
        ::"r"(tmem_d), "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc), "r"(accumulate)
        : "memory");
}

constexpr int kRowsMaxSlots = 8;     // X row ring slots (excluding 2 mirror + 1 pad slot)
constexpr int kRowsDySlots = 3;

struct WgradRowsParams {
    const __nv_bfloat16* x; int64_t csx;         // X: pixel 0 of the first plane of this launch, plane stride (elements)
    const __nv_bfloat16* dy; int64_t csd; int pd;
    float* partial;                              // [grid][3][M][N]
    int N;                                       // Cout
    int H, pw, S;
    int SW, strips, RC, chunks, units;
    int R;
    int tiles;                                   // 128-pixel tiles per row segment
    uint32_t PLB, seg_bytes, tmem_cols;
};

template <int KP>
__global__ void __launch_bounds__(192, 1) wgrad_rows_kernel(const __grid_constant__ WgradRowsParams p) {
    constexpr int M = KP == 2 ? 64 : 128;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[kRowsMaxSlots];
    __shared__ __align__(8) uint64_t empty_bar[kRowsMaxSlots];
    __shared__ __align__(8) uint64_t dfull_bar[kRowsDySlots];
    __shared__ __align__(8) uint64_t dempty_bar[kRowsDySlots];
    __shared__ __align__(8) uint64_t done_bar;
    __shared__ uint32_t s_tmem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int R = p.R;
    const uint32_t ROWB = (uint32_t)KP * p.PLB;                   // X row slot
    const uint32_t DROWB = (uint32_t)p.pd * p.PLB;                // dY row slot
    uint8_t* xring = smem;
    uint8_t* dring = smem + (uint32_t)(R + 3) * ROWB;
    const uint32_t total_bytes = (uint32_t)(R + 3) * ROWB + (uint32_t)kRowsDySlots * DROWB;

    // zero the rings once: the tails of the slots (beyond the copied segment) are read by the last tile of a segment and by
    // the padding M groups; they must hold finite values (X) / zeros (dY)
    for (uint32_t i = threadIdx.x * 16u; i < total_bytes; i += 192u * 16u) *reinterpret_cast<uint4*>(smem + i) = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (threadIdx.x == 0) {
        for (int i = 0; i < R; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < kRowsDySlots; ++i) { mbar_init(&dfull_bar[i], 1); mbar_init(&dempty_bar[i], 1); }
        mbar_init(&done_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(&s_tmem, p.tmem_cols);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = s_tmem;
    const int per_sample = p.strips * p.chunks;

    if (warp == 0) {
        if (lane == 0) {
            int xslot = 0, dslot = 0;
            uint32_t xphase = 1, dphase = 1;                      // parity to wait for on the empty barriers
            for (int unit = blockIdx.x; unit < p.units; unit += gridDim.x) {
                const int n = unit / per_sample, rem = unit - n * per_sample;
                const int strip = rem / p.chunks, chunk = rem - strip * p.chunks;
                const int ya = chunk * p.RC, rows = min(p.H, ya + p.RC) - ya;
                // padded-linear pixel of (row ya - 1, first column of the strip's segment)
                int64_t px = (int64_t)n * p.S + (int64_t)ya * p.pw + (int64_t)strip * p.SW;
                for (int i = 0; i < rows + 2; ++i, px += p.pw) {
                    mbar_wait(&empty_bar[xslot], xphase);
                    const bool mirror = xslot < 2;
                    mbar_arrive_expect_tx(&full_bar[xslot], (uint32_t)KP * p.seg_bytes * (mirror ? 2u : 1u));
#pragma unroll
                    for (int k = 0; k < KP; ++k) {
                        const __nv_bfloat16* src = p.x + (int64_t)k * p.csx + px * 8;
                        uint8_t* dst = xring + (uint32_t)xslot * ROWB + (uint32_t)k * p.PLB;
                        bulk_load_r(dst, src, p.seg_bytes, &full_bar[xslot]);
                        if (mirror) bulk_load_r(dst + (uint32_t)R * ROWB, src, p.seg_bytes, &full_bar[xslot]);
                    }
                    if (++xslot == R) { xslot = 0; xphase ^= 1u; }
                    if (i >= 2) {                                  // dY row ya + i - 2 = the row one above this X row
                        mbar_wait(&dempty_bar[dslot], dphase);
                        mbar_arrive_expect_tx(&dfull_bar[dslot], (uint32_t)p.pd * p.seg_bytes);
                        const int64_t pd_px = px - p.pw;
                        for (int k = 0; k < p.pd; ++k)
                            bulk_load_r(dring + (uint32_t)dslot * DROWB + (uint32_t)k * p.PLB, p.dy + (int64_t)k * p.csd + pd_px * 8,
                                        p.seg_bytes, &dfull_bar[dslot]);
                        if (++dslot == kRowsDySlots) { dslot = 0; dphase ^= 1u; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        const uint32_t x_base = (8u << 16) | (smem_u32(xring) >> 4);       // LBO = 128 B: next 8 pixels (K)
        const uint32_t d_base = (8u << 16) | (smem_u32(dring) >> 4);
        const uint32_t hi = (p.PLB >> 4) | (1u << 14);                     // SBO = plane stride: next 8 channels (M / N)
        const uint32_t idesc = make_idesc_bf16(M, p.N, 1, 1);
        const uint32_t ncol = (uint32_t)p.N;
        const uint32_t rowb16 = ROWB >> 4, drowb16 = DROWB >> 4;
        int rslot = 0, sslot = 0, dslot = 0;             // next X slot to wait for / slot of the current top row / dY slot
        uint32_t rphase = 0, dphase = 0;
        int ready = 0, top = 0;                          // X rows whose full barrier has been passed / rows consumed as "top"
        uint32_t acc = 0;
        for (int unit = blockIdx.x; unit < p.units; unit += gridDim.x) {
            const int rem = unit % per_sample;
            const int chunk = rem % p.chunks;
            const int ya = chunk * p.RC, rows = min(p.H, ya + p.RC) - ya;
            for (int i = 0; i < rows; ++i) {
                while (ready < top + 3) {                // rows top, top + 1, top + 2 = y - 1, y, y + 1
                    mbar_wait(&full_bar[rslot], rphase);
                    ++ready;
                    if (++rslot == R) { rslot = 0; rphase ^= 1u; }
                }
                mbar_wait(&dfull_bar[dslot], dphase);
                tc_fence_after_sync();
                if (elect_one_r()) {
                    const uint32_t xrow = x_base + (uint32_t)sslot * rowb16;
                    const uint32_t drow = d_base + (uint32_t)dslot * drowb16 + 1u;        // dY pixel = segment pixel + 1
                    for (int t = 0; t < p.tiles; ++t) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const uint32_t off = (uint32_t)(t * 128 + k * 16);
#pragma unroll
                            for (int dx = 0; dx < 3; ++dx)
                                umma_lohi_r(tmem_base + (uint32_t)dx * ncol, xrow + off + (uint32_t)dx, hi, drow + off, hi, idesc, acc);
                            acc = 1u;
                        }
                    }
                    umma_commit(&dempty_bar[dslot]);
                    umma_commit(&empty_bar[sslot]);       // the top row is below every later window of this unit
                }
                acc = 1u;
                ++top;
                if (++sslot == R) sslot = 0;
                if (++dslot == kRowsDySlots) { dslot = 0; dphase ^= 1u; }
            }
            // the unit's last two X rows are dead too (the next unit starts a fresh window)
            while (ready < top + 2) {                    // only when rows == 0 (never) - keeps the counters consistent
                mbar_wait(&full_bar[rslot], rphase);
                ++ready;
                if (++rslot == R) { rslot = 0; rphase ^= 1u; }
            }
            if (elect_one_r()) {
                umma_commit(&empty_bar[sslot]);
                umma_commit(&empty_bar[sslot + 1 == R ? 0 : sslot + 1]);
            }
            top += 2;
            sslot += 2;
            if (sslot >= R) sslot -= R;
        }
        if (elect_one_r()) umma_commit(&done_bar);
    } else {
        // epilogue: M = 64 accumulators sit in lanes 0..15 of each TMEM lane quarter (row = 16 q + lane), M = 128 in all lanes
        const int q = warp & 3;
        mbar_wait(&done_bar, 0);
        tc_fence_after_sync();
        const int row = M == 64 ? q * 16 + lane : q * 32 + lane;
        const bool has_row = M == 64 ? lane < 16 : true;
        for (int dx = 0; dx < 3; ++dx) {
            float* out = p.partial + (((int64_t)blockIdx.x * 3 + dx) * M + row) * p.N;
            for (int c0 = 0; c0 < p.N; c0 += 16) {
                float v[16];
                tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(dx * p.N + c0), v);
                if (has_row) {
#pragma unroll
                    for (int i = 0; i < 16; i += 4)
                        *reinterpret_cast<float4*>(out + c0 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, p.tmem_cols);
    }
}

// dw[co][ci_off + ci][tap] = sum over CTAs of partial[cta][dx][m][co],  m = ((dy * KP + plane) * 8 + c8), ci = plane * 8 + c8
__global__ void __launch_bounds__(256) wgrad_rows_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dw, int ctas,
                                                                 int M, int KP, int N, int Cin_total, int ci_off) {
    const int total = 9 * KP * 8 * N;
    const int i = blockIdx.x * 32 + (threadIdx.x & 31);
    const bool valid = i < total;
    const int iv = valid ? i : 0;
    const int co = iv % N, ci = (iv / N) % (KP * 8), tap = iv / (N * KP * 8);
    const int dy = tap / 3, dx = tap - dy * 3;
    const int m = (dy * KP + (ci >> 3)) * 8 + (ci & 7);
    const float t = sliced_partial_sum(partial + ((int64_t)dx * M + m) * N + co, (int64_t)3 * M * N, ctas, valid);
    if (valid && threadIdx.x < 32) dw[((int64_t)co * Cin_total + ci_off + ci) * 9 + tap] = t;
}

struct RowsPlan { int SW, strips, tiles, R, RC, chunks, units, grid, per_sm; uint32_t PLB, seg_bytes, smem_bytes, tmem_cols; bool ok; };

// One CTA per SM with 256-pixel strips and the deepest row ring, or (pass 0, when it fits) two CTAs per SM with 128-pixel strips:
// the producer / epilogue stalls of one CTA hide behind the other's MMAs (the same effect measured on conv_ring).
// MRISR_ROWS_CTAS=1 keeps the one-CTA plan.
static RowsPlan rows_plan(int N, int H, int W, int KP, int Cout) {
    static const int want = [] { const char* e = getenv("MRISR_ROWS_CTAS"); return e ? atoi(e) : 2; }();
    RowsPlan pl;
    pl.ok = false;
    const int cols = 3 * Cout;
    pl.tmem_cols = cols <= 32 ? 32 : cols <= 64 ? 64 : cols <= 128 ? 128 : cols <= 256 ? 256 : 512;
    pl.per_sm = 1;
    for (int pass = (want >= 2 && pl.tmem_cols <= 256) ? 0 : 1; pass < 2; ++pass) {
        const int sw_max = pass == 0 ? 128 : 256;
        const uint32_t limit = pass == 0 ? 113u * 1024u : 225u * 1024u;
        pl.strips = (W + sw_max - 1) / sw_max;
        if (W % pl.strips) { if (pass == 0) continue; return pl; }
        pl.SW = W / pl.strips;
        pl.tiles = (pl.SW + 127) / 128;
        pl.seg_bytes = (uint32_t)(pl.SW + 2) * 16u;
        pl.PLB = (((uint32_t)(pl.tiles * 128 + 3) * 16u) + 127u) & ~127u;
        if ((pl.PLB >> 4) >= (1u << 14)) return pl;
        const uint32_t rowb = (uint32_t)KP * pl.PLB, drowb = (uint32_t)(Cout / 8) * pl.PLB;
        pl.R = 0;
        for (int R = kRowsMaxSlots; R >= (pass == 0 ? 6 : 4); --R) {
            const uint32_t need = (uint32_t)(R + 3) * rowb + (uint32_t)kRowsDySlots * drowb + 1024u;
            if (need <= limit) { pl.R = R; pl.smem_bytes = need; break; }
        }
        if (pl.R) { pl.per_sm = pass == 0 ? 2 : 1; break; }
        if (pass == 1) return pl;
    }
    // rows per chunk: enough units to balance the CTAs (>= 6 per CTA when possible), halo overhead 2 / RC
    const int ctas = num_sms() * pl.per_sm;
    const int64_t want_units = (int64_t)ctas * 6;
    int rc = (int)(((int64_t)N * pl.strips * H) / want_units);
    if (rc < 8) rc = 8;
    if (rc > 64) rc = 64;
    if (rc > H) rc = H;
    pl.RC = rc;
    pl.chunks = (H + rc - 1) / rc;
    pl.units = N * pl.strips * pl.chunks;
    pl.grid = pl.units < ctas ? pl.units : ctas;
    pl.ok = true;
    return pl;
}

}  // namespace tc

using bf16 = __nv_bfloat16;

// plane groups a source of C channels is processed in: 4 planes at a time, a 16-channel tensor as one group of 2
static inline int rows_group_planes(int C) { return C == 16 ? 2 : 4; }

bool wgrad_rows_supported(int N, int H, int W, int C1, int C2, int Cout, int ks) {
    if (ks != 3 || !(Cout == 16 || Cout == 32 || Cout == 64)) return false;
    if (!(C1 == 16 || C1 == 32 || C1 == 64) || !(C2 == 0 || C2 == 16 || C2 == 32 || C2 == 64)) return false;
    if (H < 2 || W < 8) return false;
    return tc::rows_plan(N, H, W, rows_group_planes(C1), Cout).ok && (C2 == 0 || tc::rows_plan(N, H, W, rows_group_planes(C2), Cout).ok);
}

size_t wgrad_rows_scratch_elems(int N, int H, int W, int Cin, int Cout) {
    return (size_t)tc::num_sms() * 2 * 3 * 128 * Cout;          // up to two CTAs per SM, one partial each
}

static int wgrad_rows_launch_group(const bf16* x, int64_t csx, int KP, const WgradArgs<bf16>& a, int Cin_total, int ci_off, cudaStream_t s) {
    const int H = a.dy.H, W = a.dy.W, N = a.dy.N, Cout = a.dy.C;
    const tc::RowsPlan pl = tc::rows_plan(N, H, W, KP, Cout);
    tc::WgradRowsParams p;
    p.x = x; p.csx = csx;
    p.dy = a.dy.p; p.csd = a.dy.cs; p.pd = Cout / 8;
    p.partial = a.partial; p.N = Cout;
    p.H = H; p.pw = W + 2; p.S = (H + 2) * (W + 2);
    p.SW = pl.SW; p.strips = pl.strips; p.RC = pl.RC; p.chunks = pl.chunks; p.units = pl.units;
    p.R = pl.R; p.tiles = pl.tiles; p.PLB = pl.PLB; p.seg_bytes = pl.seg_bytes; p.tmem_cols = pl.tmem_cols;
    static PerDeviceFlag attr_pd[2]; bool attr_done_v[2] = {attr_pd[0].cur(), attr_pd[1].cur()}; bool* attr_done = attr_done_v;
    if (KP == 2) {
        if (!attr_done[0]) {
            MRISR_CUDA_CHECK(cudaFuncSetAttribute(tc::wgrad_rows_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
            attr_pd[0].cur() = true;
        }
        tc::wgrad_rows_kernel<2><<<pl.grid, 192, pl.smem_bytes, s>>>(p);
    } else {
        if (!attr_done[1]) {
            MRISR_CUDA_CHECK(cudaFuncSetAttribute(tc::wgrad_rows_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
            attr_pd[1].cur() = true;
        }
        tc::wgrad_rows_kernel<4><<<pl.grid, 192, pl.smem_bytes, s>>>(p);
    }
    MRISR_LAUNCH_CHECK();
    const int M = KP == 2 ? 64 : 128;
    const int total = 9 * KP * 8 * Cout;
    tc::wgrad_rows_reduce_kernel<<<ceil_div(total, 32), 256, 0, s>>>(a.partial, a.dw, pl.grid, M, KP, Cout, Cin_total, ci_off);
    MRISR_LAUNCH_CHECK();
    return 0;
}

int launch_wgrad_rows(const WgradArgs<bf16>& a, cudaStream_t s) {
    const int C1 = a.src1.C, C2 = a.src2.p ? a.src2.C : 0, Cin = C1 + C2, Cout = a.dy.C;
    const int H = a.dy.H, W = a.dy.W, N = a.dy.N;
    if (!wgrad_rows_supported(N, H, W, C1, C2, Cout, a.ks) || !a.src1.planar() || (C2 && !a.src2.planar()) || !a.dy.planar()) {
        set_error("wgrad_rows: unsupported shape/layout C1=%d C2=%d Cout=%d ks=%d", C1, C2, Cout, a.ks);
        return -2;
    }
    const char* tag = "wgrad_rows";
    char tbuf[96];
    if (prof_on()) {
        snprintf(tbuf, sizeof(tbuf), "wgrad_rows:k%d:%d>%d@%dx%d", a.ks, Cin, Cout, H, W);
        tag = prof_intern(tbuf);
    }
    ProfScope prof_(tag, s);
    int ci_off = 0;
    for (int srcn = 0; srcn < 2; ++srcn) {
        const Act<bf16>& t = srcn == 0 ? a.src1 : a.src2;
        if (srcn == 1 && !C2) break;
        const int gp = rows_group_planes(t.C);
        for (int k0 = 0; k0 < t.C / 8; k0 += gp) {
            int rc = wgrad_rows_launch_group(t.p + (int64_t)k0 * t.cs, t.cs, gp, a, Cin, ci_off, s);
            if (rc) return rc;
            ci_off += gp * 8;
        }
    }
    return 0;
}

}  // namespace mrisr
