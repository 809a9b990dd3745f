// tcgen05 weight gradient for the narrow high-resolution layers over the planar-8 layout (companion of conv_ring.cu):
//   dW[tap][ci][co] = sum_p X[p + shift(tap)][ci] * dY[p][co]           (autograd of nn.Conv2d, models/unet_model.py:29..168)
// The reduction runs over pixels, so both operands are "MN-major": 8 channels contiguous, pixels along K - which is
// exactly a planar-8 block in shared memory (no-swizzle canonical layout: channel groups SBO = plane stride apart, groups of
// 8 pixels LBO = 128 B apart).  A CTA walks a contiguous run of 128-pixel blocks with the same sliding ring of X blocks as
// conv_ring.cu (every input byte fetched once, a tap = descriptor start moved by shift * 16 B) plus a small ring of dY
// blocks, and accumulates one 64 x N fp32 tile per tap in TMEM (M = 64: the 16..64 channels of one operand padded to 64
// rows; N = the other operand's channels, <= 32).  Per-CTA partials are reduced in a fixed order by
// wgrad_ring_reduce_kernel into the reference's [Cout][Cin][kh][kw] layout.
#include "kernels.cuh"
#include "tc_common.cuh"

namespace mrisr {
namespace tc {

int num_sms();

__device__ __forceinline__ void bulk_load_w(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}


__device__ __forceinline__ bool elect_one_w() {
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
// tcgen05.mma with both descriptors given as (lo, hi) halves
__device__ __forceinline__ void umma_bf16_lohi2(uint32_t tmem_d, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi,
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

constexpr int kWRingMaxSlots = 16;
constexpr int kDySlots = 3;

struct WgradRingParams {
    const __nv_bfloat16* x1; int64_t csx1; int px1;     // shifted operand: forward input (1 or 2 planar sources)
    const __nv_bfloat16* x2; int64_t csx2; int px2;
    const __nv_bfloat16* dy; int64_t csd; int pd;       // gradient wrt the raw conv output (zero border)
    float* partial;                                     // [CTAs][taps][64][N]
    int a_is_x;                                         // 1: M side = X channels (exactly 64), N = Cout;  0: M side = dY, N = Cin
    int N;                                              // GEMM N
    int taps, pw, num_tiles, tiles_per_cta;
    int R, halo_lo, halo_hi;
    int sh[9];                                          // halo_lo * 128 + pixel shift of tap t (>= 0)
    uint32_t tmem_cols;
};

__global__ void __launch_bounds__(192, 1) wgrad_ring_kernel(const __grid_constant__ WgradRingParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[kWRingMaxSlots];
    __shared__ __align__(8) uint64_t empty_bar[kWRingMaxSlots];
    __shared__ __align__(8) uint64_t dfull_bar[kDySlots];
    __shared__ __align__(8) uint64_t dempty_bar[kDySlots];
    __shared__ __align__(8) uint64_t done_bar;
    __shared__ uint32_t s_tmem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KPX = p.px1 + p.px2;
    const uint32_t PLSX = (uint32_t)(p.R + 1) * 2048u;            // plane stride of the X ring (incl. mirror slot)
    const uint32_t PLSD = (uint32_t)kDySlots * 2048u;             // plane stride of the dY ring
    // the M-side operand is read as 8 channel groups whatever its real width: keep it FIRST so the padding groups of a
    // narrow dY fall inside this CTA's shared memory (they only produce rows that are never stored)
    uint8_t* dring = p.a_is_x ? smem + (uint32_t)KPX * PLSX : smem;
    uint8_t* xring = p.a_is_x ? smem : smem + (uint32_t)p.pd * PLSD;
    const int t_begin = blockIdx.x * p.tiles_per_cta;
    const int t_end = min(p.num_tiles, t_begin + p.tiles_per_cta);
    const int b0 = t_begin - p.halo_lo;
    const int num_blocks = (t_end - 1 + p.halo_hi) - b0 + 1;

    if (threadIdx.x == 0) {
        for (int i = 0; i < p.R; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < kDySlots; ++i) { mbar_init(&dfull_bar[i], 1); mbar_init(&dempty_bar[i], 1); }
        mbar_init(&done_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(&s_tmem, p.tmem_cols);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = s_tmem;

    if (warp == 0) {
        if (lane == 0 && t_begin < t_end) {
            // X block j and dY block of tile (t_begin + jd) are issued interleaved so neither ring starves the other
            int j = 0, jd = 0;
            const int n_tiles = t_end - t_begin;
            while (j < num_blocks || jd < n_tiles) {
                // dY block jd is only useful once the X window of its tile has been requested
                const bool want_x = j < num_blocks && (jd >= n_tiles || j <= jd + p.halo_lo + p.halo_hi);
                if (want_x) {
                    const int slot = j % p.R;
                    mbar_wait(&empty_bar[slot], ((j / p.R) & 1) ^ 1);
                    const bool mirror = slot == 0;
                    mbar_arrive_expect_tx(&full_bar[slot], (uint32_t)KPX * 2048u * (mirror ? 2u : 1u));
                    const int64_t pix0 = (int64_t)(b0 + j) * 128;
                    for (int k = 0; k < KPX; ++k) {
                        const __nv_bfloat16* src = k < p.px1 ? p.x1 + (int64_t)k * p.csx1 + pix0 * 8
                                                             : p.x2 + (int64_t)(k - p.px1) * p.csx2 + pix0 * 8;
                        uint8_t* dst = xring + (uint32_t)k * PLSX + (uint32_t)slot * 2048u;
                        bulk_load_w(dst, src, 2048u, &full_bar[slot]);
                        if (mirror) bulk_load_w(dst + (uint32_t)p.R * 2048u, src, 2048u, &full_bar[slot]);
                    }
                    ++j;
                } else {
                    const int slot = jd % kDySlots;
                    mbar_wait(&dempty_bar[slot], ((jd / kDySlots) & 1) ^ 1);
                    mbar_arrive_expect_tx(&dfull_bar[slot], (uint32_t)p.pd * 2048u);
                    const int64_t pix0 = (int64_t)(t_begin + jd) * 128;
                    for (int k = 0; k < p.pd; ++k)
                        bulk_load_w(dring + (uint32_t)k * PLSD + (uint32_t)slot * 2048u, p.dy + (int64_t)k * p.csd + pix0 * 8, 2048u,
                                    &dfull_bar[slot]);
                    ++jd;
                }
            }
        }
    } else if (warp == 1) {
        // warp-uniform issue loop, fully unrolled over (K step, tap): a descriptor is one integer add (the mirror slot
        // covers the 112-pixel reach of the K steps, so only the per-tile tap offsets need the ring wrap)
        if (t_begin < t_end) {
            const uint32_t ring_px = (uint32_t)p.R * 128u;
            const uint32_t x_base = (8u << 16) | (smem_u32(xring) >> 4);          // LBO = 128 B (next 8 pixels)
            const uint32_t d_base = (8u << 16) | (smem_u32(dring) >> 4);
            const uint32_t x_hi = (PLSX >> 4) | (1u << 14);                       // SBO = plane stride (next 8 channels)
            const uint32_t d_hi = (PLSD >> 4) | (1u << 14);
            const uint32_t idesc = make_idesc_bf16(64, p.N, 1, 1);
            const uint32_t ncol = (uint32_t)p.N;
            const bool a_is_x = p.a_is_x != 0;
            uint32_t sh[9];
#pragma unroll
            for (int t = 0; t < 9; ++t) sh[t] = (uint32_t)p.sh[t];
            const int first_need = p.halo_lo + p.halo_hi;
            int ready = 0, rslot = 0, eslot = 0, dslot = 0;
            uint32_t rphase = 0, dphase = 0, tb = 0;
            const int n_tiles = t_end - t_begin;
            for (int it = 0; it < n_tiles; ++it) {
                const int need = it + first_need;
                while (ready <= need) {
                    mbar_wait(&full_bar[rslot], rphase);
                    ++ready;
                    if (++rslot == p.R) { rslot = 0; rphase ^= 1u; }
                }
                mbar_wait(&dfull_bar[dslot], dphase);
                tc_fence_after_sync();
                const bool leader = elect_one_w();
                uint32_t off[9];
#pragma unroll
                for (int t = 0; t < 9; ++t) {
                    off[t] = tb + sh[t];
                    if (off[t] >= ring_px) off[t] -= ring_px;
                }
                const uint32_t dlo0 = d_base + (uint32_t)dslot * 128u;
                if (!leader) {
                } else if (p.taps == 9) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) {                          // 8 K steps of 16 pixels
#pragma unroll
                        for (int tap = 0; tap < 9; ++tap) {
                            const uint32_t xlo = x_base + off[tap] + (uint32_t)(k * 16);
                            const uint32_t dlo = dlo0 + (uint32_t)(k * 16);
                            const uint32_t acc = (it > 0 || k > 0) ? 1u : 0u;
                            if (a_is_x) umma_bf16_lohi2(tmem_base + (uint32_t)tap * ncol, xlo, x_hi, dlo, d_hi, idesc, acc);
                            else umma_bf16_lohi2(tmem_base + (uint32_t)tap * ncol, dlo, d_hi, xlo, x_hi, idesc, acc);
                        }
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const uint32_t xlo = x_base + off[0] + (uint32_t)(k * 16);
                        const uint32_t dlo = dlo0 + (uint32_t)(k * 16);
                        const uint32_t acc = (it > 0 || k > 0) ? 1u : 0u;
                        if (a_is_x) umma_bf16_lohi2(tmem_base, xlo, x_hi, dlo, d_hi, idesc, acc);
                        else umma_bf16_lohi2(tmem_base, dlo, d_hi, xlo, x_hi, idesc, acc);
                    }
                }
                if (leader) {
                    umma_commit(&dempty_bar[dslot]);
                    umma_commit(&empty_bar[eslot]);                       // X block `it` is below every later window
                }
                if (++dslot == kDySlots) { dslot = 0; dphase ^= 1u; }
                if (++eslot == p.R) eslot = 0;
                tb += 128u;
                if (tb >= ring_px) tb -= ring_px;
            }
            if (elect_one_w()) umma_commit(&done_bar);
        }
    } else if (t_begin < t_end) {
        // epilogue: M = 64 accumulators occupy lanes 0..15 of each TMEM lane quarter (row = 16 * quarter + lane)
        const int q = warp & 3;
        mbar_wait(&done_bar, 0);
        tc_fence_after_sync();
        const int row = q * 16 + lane;
        for (int tap = 0; tap < p.taps; ++tap) {
            float* out = p.partial + (((int64_t)blockIdx.x * p.taps + tap) * 64 + row) * p.N;
            for (int c0 = 0; c0 < p.N; c0 += 16) {
                float v[16];
                tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(tap * p.N + c0), v);
                if (lane < 16) {
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

// dw[co][ci][tap] = sum over CTAs of partial[cta][tap][m][n];  (m, n) = (ci, co) if a_is_x else (co, ci)
__global__ void wgrad_ring_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dw, int ctas, int taps, int N,
                                         int Cin, int Cout, int a_is_x) {
    const int total = taps * Cin * Cout;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int tap = i / (Cin * Cout), r = i - tap * Cin * Cout;
    int m, n;
    if (a_is_x) { m = r / Cout; n = r - m * Cout; } else { m = r / Cin; n = r - m * Cin; }
    float t = 0.f;
    for (int c = 0; c < ctas; ++c) t += partial[(((int64_t)c * taps + tap) * 64 + m) * N + n];
    const int ci = a_is_x ? m : n, co = a_is_x ? n : m;
    dw[((int64_t)co * Cin + ci) * taps + tap] = t;
}

struct WRingPlan { int R, halo_lo, halo_hi, a_is_x, N; uint32_t smem_bytes, tmem_cols; bool ok; };

static WRingPlan wring_plan(int W, int Cin, int Cout, int ks) {
    WRingPlan pl;
    const int pw = W + 2;
    pl.halo_lo = ks == 3 ? (pw + 1 + 127) / 128 : 0;
    pl.halo_hi = ks == 3 ? (pw + 128 + 15) / 128 : 0;           // K steps start up to 112 pixels into the block
    const int live = pl.halo_lo + pl.halo_hi + 1;
    pl.a_is_x = Cin == 64 ? 1 : 0;
    pl.N = pl.a_is_x ? Cout : Cin;
    const int cols = ks * ks * pl.N;
    pl.tmem_cols = cols <= 32 ? 32 : cols <= 64 ? 64 : cols <= 128 ? 128 : cols <= 256 ? 256 : 512;
    pl.ok = false; pl.R = 0; pl.smem_bytes = 0;
    if (cols > 512 || pl.N > 32 || pl.N < 16) return pl;
    for (int ahead = 3; ahead >= 1; --ahead) {
        const int R = live + ahead;
        uint32_t need = (uint32_t)(Cin / 8) * (uint32_t)(R + 1) * 2048u + (uint32_t)(Cout / 8) * kDySlots * 2048u + 1024u;
        const uint32_t pad_reach = 8u * kDySlots * 2048u + 4096u;       // reach of the padded M groups of a narrow dY
        if (need < pad_reach) need = pad_reach;
        if (R <= kWRingMaxSlots && need <= 200u * 1024u) { pl.R = R; pl.smem_bytes = need; pl.ok = true; break; }
    }
    return pl;
}

}  // namespace tc

using bf16 = __nv_bfloat16;

bool wgrad_ring_supported(int H, int W, int C1, int C2, int Cout, int ks) {
    const int Cin = C1 + C2;
    if (!(Cout == 16 || Cout == 32 || Cout == 64)) return false;
    if (C1 < 8 || C1 % 8 || C2 % 8 || Cin % 16 || Cin > 64 || (ks != 1 && ks != 3)) return false;
    return tc::wring_plan(W, Cin, Cout, ks).ok;
}

size_t wgrad_ring_scratch_elems(int N, int H, int W, int Cin, int Cout, int ks) {
    const tc::WRingPlan pl = tc::wring_plan(W, Cin, Cout, ks);
    return (size_t)tc::num_sms() * ks * ks * 64 * pl.N;
}

int launch_wgrad_ring(const WgradArgs<bf16>& a, cudaStream_t s) {
    const int C1 = a.src1.C, C2 = a.src2.p ? a.src2.C : 0, Cin = C1 + C2, Cout = a.dy.C;
    const int H = a.dy.H, W = a.dy.W, N = a.dy.N;
    if (!wgrad_ring_supported(H, W, C1, C2, Cout, a.ks) || !a.src1.planar() || (C2 && !a.src2.planar()) || !a.dy.planar()) {
        set_error("wgrad_ring: unsupported shape/layout C1=%d C2=%d Cout=%d ks=%d", C1, C2, Cout, a.ks);
        return -2;
    }
    const char* tag = "wgrad_ring";
    char tbuf[96];
    if (prof_on()) {
        snprintf(tbuf, sizeof(tbuf), "wgrad_ring:k%d:%d>%d@%dx%d", a.ks, Cin, Cout, H, W);
        tag = prof_intern(tbuf);
    }
    ProfScope prof_(tag, s);
    const tc::WRingPlan pl = tc::wring_plan(W, Cin, Cout, a.ks);
    tc::WgradRingParams p;
    p.x1 = a.src1.p; p.csx1 = a.src1.cs; p.px1 = C1 / 8;
    p.x2 = C2 ? a.src2.p : a.src1.p; p.csx2 = C2 ? a.src2.cs : 0; p.px2 = C2 / 8;
    p.dy = a.dy.p; p.csd = a.dy.cs; p.pd = Cout / 8;
    p.partial = a.partial; p.a_is_x = pl.a_is_x; p.N = pl.N;
    p.taps = a.ks * a.ks; p.pw = W + 2;
    const int64_t rows = (int64_t)N * (H + 2) * (W + 2);
    p.num_tiles = (int)((rows + 127) / 128);
    p.tiles_per_cta = ceil_div(p.num_tiles, tc::num_sms());
    p.R = pl.R; p.halo_lo = pl.halo_lo; p.halo_hi = pl.halo_hi; p.tmem_cols = pl.tmem_cols;
    for (int t = 0; t < 9; ++t) p.sh[t] = pl.halo_lo * 128 + (p.taps == 9 ? (t / 3 - 1) * p.pw + (t % 3 - 1) : 0);
    const int grid = ceil_div(p.num_tiles, p.tiles_per_cta);
    static PerDeviceFlag attr_done_pd; bool& attr_done = attr_done_pd.cur();
    if (!attr_done) {
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(tc::wgrad_ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 201 * 1024));
        attr_done = true;
    }
    tc::wgrad_ring_kernel<<<grid, 192, pl.smem_bytes, s>>>(p);
    MRISR_LAUNCH_CHECK();
    const int total = p.taps * Cin * Cout;
    tc::wgrad_ring_reduce_kernel<<<ceil_div(total, 256), 256, 0, s>>>(a.partial, a.dw, grid, p.taps, pl.N, Cin, Cout, pl.a_is_x);
    MRISR_LAUNCH_CHECK();
    return 0;
}

}  // namespace mrisr
