// tcgen05 weight-gradient kernel for 3x3 / 1x1 convolutions over padded NHWC bf16 activations (fp32 accumulate in TMEM).
//   dW[tap][ci][co] = sum_rows X[row + shift(tap)][ci] * dY[row][co]        (autograd of nn.Conv2d, models/unet_model.py:29..172)
// where `row` runs over the padded-linear pixel index of conv_tc.cu; dY has a zero border, so border rows add nothing and
// no masking is needed.  As a GEMM:  D[(tap,ci) x co] = A * B with the reduction over pixels, i.e. both operands are read
// "MN-major" (channels contiguous, pixels along K) exactly as TMA delivers a [64 pixels x CS channels] box of an NHWC
// tensor with a CS*2-byte swizzle.  M = 128 rows = 128/CS slabs, each slab one (tap, channel-chunk) box of X shifted by the
// tap offset; N = Cout (Cout/64 slabs of dY).  A CTA owns MT M-tiles (MT*Cout <= 512 TMEM columns) and one split of the
// pixel range; split partials are summed, in a fixed order, by wgrad_tc_reduce_kernel into the reference's
// [Cout][Cin][kh][kw] fp32 layout.
#include <stdlib.h>
#include "kernels.cuh"
#include "tc_common.cuh"

namespace mrisr {
namespace tc {

int num_sms();


__device__ __forceinline__ bool elect_one_t() {
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
__device__ __forceinline__ void umma_lohi_t(uint32_t tmem_d, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi,
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

constexpr int kBKP = 64;          // pixels (GEMM K) per pipeline stage

struct WgradTcParams {
    float* partial;               // [splits][Mtot][BN]
    int BN, CS, CSB, MT, stages;
    int taps, pw, chunks1, chunks2, num_slabs, Mtot;
    int num_kb, kb_per_split;
    uint32_t a_tile_bytes, a_slab_bytes, b_slab_bytes, b_bytes, stage_bytes;
    uint32_t lbo_a, sbo_a, lt_a, lbo_b, sbo_b, lt_b, idesc, tmem_cols;
};

constexpr int kMaxStages = 8;

__global__ void __launch_bounds__(192, 1)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmX1, const __grid_constant__ CUtensorMap tmX2,
                const __grid_constant__ CUtensorMap tmDY, const WgradTcParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[kMaxStages];
    __shared__ __align__(8) uint64_t empty_bar[kMaxStages];
    __shared__ __align__(8) uint64_t done_bar;
    __shared__ uint32_t s_tmem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int SPT = 128 / p.CS;                                   // slabs per M-tile
    const int cpt = p.chunks1 + p.chunks2;                        // slabs per tap
    const int tile0 = blockIdx.x * p.MT;                          // first M-tile of this CTA
    const int kb_begin = blockIdx.y * p.kb_per_split;
    const int kb_end = min(p.num_kb, kb_begin + p.kb_per_split);

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmX1);
        tma_prefetch_desc(&tmX2);
        tma_prefetch_desc(&tmDY);
        for (int i = 0; i < p.stages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        mbar_init(&done_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(&s_tmem, p.tmem_cols);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = s_tmem;

    if (warp == 0) {
        if (lane == 0) {
            // bytes landing per stage: the valid X slabs of this CTA + the whole dY block
            int valid_slabs = 0;
            for (int j = 0; j < p.MT; ++j)
                for (int i = 0; i < SPT; ++i)
                    if ((tile0 + j) * SPT + i < p.num_slabs) ++valid_slabs;
            const uint32_t tx_bytes = (uint32_t)valid_slabs * p.a_slab_bytes + p.b_bytes;
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = kb_begin; kb < kb_end; ++kb) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
                uint8_t* base = smem + (size_t)stage * p.stage_bytes;
                const int row0 = kb * kBKP;
                for (int j = 0; j < p.MT; ++j) {
                    for (int i = 0; i < SPT; ++i) {
                        const int sidx = (tile0 + j) * SPT + i;
                        if (sidx >= p.num_slabs) continue;
                        const int tap = sidx / cpt, chunk = sidx - tap * cpt;
                        const int shift = p.taps == 9 ? (tap / 3 - 1) * p.pw + (tap % 3 - 1) : 0;
                        uint8_t* dst = base + (size_t)j * p.a_tile_bytes + (size_t)i * p.a_slab_bytes;
                        if (chunk < p.chunks1) tma_load_2d(dst, &tmX1, &full_bar[stage], chunk * p.CS, row0 + shift);
                        else tma_load_2d(dst, &tmX2, &full_bar[stage], (chunk - p.chunks1) * p.CS, row0 + shift);
                    }
                }
                uint8_t* bdst = base + (size_t)p.MT * p.a_tile_bytes;
                for (int b = 0; b < p.BN / p.CSB; ++b)
                    tma_load_2d(bdst + (size_t)b * p.b_slab_bytes, &tmDY, &full_bar[stage], b * p.CSB, row0);
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // warp-uniform issue loop; a descriptor is (constant hi word, lo word = base + one add)
        const uint32_t a_hi = ((p.sbo_a >> 4) & 0x3FFFu) | (1u << 14) | ((p.lt_a & 7u) << 29);
        const uint32_t b_hi = ((p.sbo_b >> 4) & 0x3FFFu) | (1u << 14) | ((p.lt_b & 7u) << 29);
        const uint32_t a_lo0 = (((p.lbo_a >> 4) & 0x3FFFu) << 16) | (smem_u32(smem) >> 4);
        const uint32_t b_lo0 = (((p.lbo_b >> 4) & 0x3FFFu) << 16) | ((smem_u32(smem) + (uint32_t)p.MT * p.a_tile_bytes) >> 4);
        const uint32_t stage16 = p.stage_bytes >> 4, tile16 = p.a_tile_bytes >> 4;
        const uint32_t ka16 = (2u * p.sbo_a) >> 4, kb16 = (2u * p.sbo_b) >> 4;
        int mt_valid = 0;
        for (int j = 0; j < p.MT; ++j) if ((tile0 + j) * SPT < p.num_slabs) ++mt_valid;
        const uint32_t bn = (uint32_t)p.BN, idesc = p.idesc;
        int stage = 0;
        uint32_t phase = 0, acc = 0;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after_sync();
            if (elect_one_t()) {
                const uint32_t a0 = a_lo0 + (uint32_t)stage * stage16, b0 = b_lo0 + (uint32_t)stage * stage16;
#pragma unroll
                for (int k = 0; k < kBKP / 16; ++k) {
                    const uint32_t blo = b0 + (uint32_t)k * kb16;
                    for (int j = 0; j < mt_valid; ++j)
                        umma_lohi_t(tmem_base + (uint32_t)j * bn, a0 + (uint32_t)j * tile16 + (uint32_t)k * ka16, a_hi, blo, b_hi, idesc, acc);
                    acc = 1u;
                }
                umma_commit(&empty_bar[stage]);
            }
            acc = 1u;
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        if (elect_one_t()) umma_commit(&done_bar);
    } else {
        const int q = warp & 3;
        mbar_wait(&done_bar, 0);
        tc_fence_after_sync();
        for (int j = 0; j < p.MT; ++j) {
            const int rowM = (tile0 + j) * 128 + q * 32 + lane;
            float* out = p.partial + ((int64_t)blockIdx.y * p.Mtot + rowM) * p.BN;
            for (int c0 = 0; c0 < p.BN; c0 += 16) {
                float v[16];
                tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(j * p.BN + c0), v);
                if (rowM < p.Mtot) {
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

// dw[co][ci][tap] = sum_s partial[s][tap * Cin + ci][co]   (fixed order -> deterministic)
__global__ void __launch_bounds__(256) wgrad_tc_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dw, int Mtot,
                                                               int Cout, int Cin, int taps, int splits) {
    const int64_t total = (int64_t)Mtot * Cout;
    const int64_t i = (int64_t)blockIdx.x * 32 + (threadIdx.x & 31);
    const bool valid = i < total;
    const float t = sliced_partial_sum(partial + (valid ? i : 0), total, splits, valid);
    if (!valid || threadIdx.x >= 32) return;
    const int co = (int)(i % Cout), m = (int)(i / Cout);
    const int tap = m / Cin, ci = m - tap * Cin;
    dw[((int64_t)co * Cin + ci) * taps + tap] = t;
}

// few splits: one thread per output, serial over the splits
__global__ void wgrad_tc_reduce_serial_kernel(const float* __restrict__ partial, float* __restrict__ dw, int Mtot, int Cout, int Cin,
                                              int taps, int splits) {
    const int64_t total = (int64_t)Mtot * Cout;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    float t = 0.f;
    for (int s = 0; s < splits; ++s) t += partial[(int64_t)s * total + i];
    const int co = (int)(i % Cout), m = (int)(i / Cout);
    const int tap = m / Cin, ci = m - tap * Cin;
    dw[((int64_t)co * Cin + ci) * taps + tap] = t;
}

struct WgradPlan {
    int CS, CSB, MT, stages, num_slabs, Mtiles, groups, splits, kb_per_split, num_kb;
    uint32_t stage_bytes, smem_bytes;
};

static WgradPlan make_plan(int N, int H, int W, int C1, int C2, int Cout, int ks) {
    WgradPlan pl;
    pl.CS = (C1 % 64 == 0 && C2 % 64 == 0) ? 64 : (C1 % 32 == 0 && C2 % 32 == 0) ? 32 : 16;
    pl.CSB = Cout < 64 ? Cout : 64;
    const int taps = ks * ks;
    pl.num_slabs = taps * ((C1 + C2) / pl.CS);
    const int SPT = 128 / pl.CS;
    pl.Mtiles = ceil_div(pl.num_slabs, SPT);
    // M-tiles per CTA: as many as TMEM holds (MT * Cout <= 512 columns) and one 2-stage pipeline fits, then spread the
    // M-tiles EVENLY over the CTA groups (an uneven last group idles its share of the SMs: all groups get the same K split)
    int mt_max = 512 / Cout;
    while (mt_max > 1 && 2u * ((uint32_t)mt_max * 16384u + (uint32_t)Cout * kBKP * 2u) > 200u * 1024u) --mt_max;
    pl.groups = ceil_div(pl.Mtiles, mt_max);
    pl.MT = ceil_div(pl.Mtiles, pl.groups);
    pl.stage_bytes = (uint32_t)pl.MT * 16384u + (uint32_t)Cout * kBKP * 2u;
    int st = (int)((200u * 1024u) / pl.stage_bytes);
    pl.stages = st > kMaxStages ? kMaxStages : st;
    pl.smem_bytes = (uint32_t)pl.stages * pl.stage_bytes + 1024u;
    const int64_t rows = (int64_t)N * (H + 2) * (W + 2);
    pl.num_kb = (int)ceil_div64(rows, kBKP);
    int sp = num_sms() / pl.groups;
    if (sp < 1) sp = 1;
    if (sp > pl.num_kb) sp = pl.num_kb;
    pl.kb_per_split = ceil_div(pl.num_kb, sp);
    pl.splits = ceil_div(pl.num_kb, pl.kb_per_split);
    return pl;
}

}  // namespace tc

using bf16 = __nv_bfloat16;

bool wgrad_tc_supported(int H, int W, int C1, int C2, int Cout, int ks) {
    // Cout = 512 (the bottleneck of base_filters = 64) runs as two 256-column halves of dY (launch_wgrad_tc)
    const bool cout_ok = Cout == 16 || Cout == 32 || Cout == 64 || Cout == 128 || Cout == 256 || Cout == 512;
    return cout_ok && C1 >= 16 && C1 % 16 == 0 && C2 % 16 == 0 && (ks == 1 || ks == 3);
}

size_t wgrad_tc_scratch_elems(int N, int H, int W, int C1, int C2, int Cout, int ks) {
    if (Cout > 256) Cout = 256;                                  // the halves run one after the other through the same scratch
    const tc::WgradPlan pl = tc::make_plan(N, H, W, C1, C2, Cout, ks);
    const size_t own = (size_t)pl.splits * (size_t)ks * ks * (C1 + C2) * Cout;
    const size_t merged = wgrad_tc3_supported(N, H, W, C1, C2, Cout, ks) ? wgrad_tc3_scratch_elems(N, H, W, C1, C2, Cout) : 0;
    return own > merged ? own : merged;
}

// MRISR_WGRAD_TC3=0 keeps the one-box-per-tap kernel for every shape (A/B measurements)
static bool use_tc3() {
    static const bool on = [] { const char* e = getenv("MRISR_WGRAD_TC3"); return !(e && e[0] == '0'); }();
    return on;
}

int launch_wgrad_tc(const WgradArgs<bf16>& a, cudaStream_t s) {
    const int C1 = a.src1.C, C2 = a.src2.p ? a.src2.C : 0, Cin = C1 + C2, Cout = a.dy.C;
    const int H = a.dy.H, W = a.dy.W, N = a.dy.N;
    if (!wgrad_tc_supported(H, W, C1, C2, Cout, a.ks)) {
        set_error("wgrad_tc: unsupported shape C1=%d C2=%d Cout=%d ks=%d", C1, C2, Cout, a.ks);
        return -2;
    }
    if (Cout == 512) {
        // one MMA holds N <= 256: two launches over the channel halves of dY (same pixel stride, dw rows [256 h, 256 h + 256))
        for (int h = 0; h < 2; ++h) {
            WgradArgs<bf16> b = a;
            b.dy.p = a.dy.p + 256 * h;
            b.dy.C = 256;
            b.dy.ps = a.dy.ps ? a.dy.ps : a.dy.C;
            b.dw = a.dw + (size_t)256 * h * Cin * a.ks * a.ks;
            const int rc2 = launch_wgrad_tc(b, s);
            if (rc2) return rc2;
        }
        return 0;
    }
    if (use_tc3() && wgrad_tc3_supported(N, H, W, C1, C2, Cout, a.ks)) return launch_wgrad_tc3(a, s);
    const char* tag = "wgrad_tc";
    char tbuf[96];
    if (prof_on()) {
        snprintf(tbuf, sizeof(tbuf), "wgrad_tc:k%d:%d>%d@%dx%d", a.ks, Cin, Cout, H, W);
        tag = prof_intern(tbuf);
    }
    ProfScope prof_(tag, s);
    const tc::WgradPlan pl = tc::make_plan(N, H, W, C1, C2, Cout, a.ks);
    const int64_t rows = (int64_t)N * (H + 2) * (W + 2);
    CUtensorMap tX1, tX2, tDY;
    int rc = tc::make_tmap_2d_bf16(&tX1, a.src1.p, (uint64_t)rows, (uint64_t)C1, tc::kBKP, (uint32_t)pl.CS);
    if (rc) return rc;
    if (C2) { rc = tc::make_tmap_2d_bf16(&tX2, a.src2.p, (uint64_t)rows, (uint64_t)C2, tc::kBKP, (uint32_t)pl.CS); if (rc) return rc; }
    else tX2 = tX1;
    rc = tc::make_tmap_2d_bf16(&tDY, a.dy.p, (uint64_t)rows, (uint64_t)Cout, tc::kBKP, (uint32_t)pl.CSB,
                               (uint64_t)(a.dy.ps > Cout ? a.dy.ps : 0));
    if (rc) return rc;

    tc::WgradTcParams p;
    p.partial = a.partial;
    p.BN = Cout; p.CS = pl.CS; p.CSB = pl.CSB; p.MT = pl.MT; p.stages = pl.stages;
    p.taps = a.ks * a.ks; p.pw = W + 2; p.chunks1 = C1 / pl.CS; p.chunks2 = C2 / pl.CS;
    p.num_slabs = pl.num_slabs; p.Mtot = p.taps * Cin;
    p.num_kb = pl.num_kb; p.kb_per_split = pl.kb_per_split;
    p.a_tile_bytes = 16384u;
    p.a_slab_bytes = (uint32_t)tc::kBKP * pl.CS * 2u;
    p.b_slab_bytes = (uint32_t)tc::kBKP * pl.CSB * 2u;
    p.b_bytes = (uint32_t)tc::kBKP * Cout * 2u;
    p.stage_bytes = pl.stage_bytes;
    p.lbo_a = p.a_slab_bytes; p.sbo_a = 8u * pl.CS * 2u; p.lt_a = tc::layout_type_for(pl.CS * 2);
    p.lbo_b = p.b_slab_bytes; p.sbo_b = 8u * pl.CSB * 2u; p.lt_b = tc::layout_type_for(pl.CSB * 2);
    p.idesc = tc::make_idesc_bf16(128, Cout, 1, 1);
    const int cols = pl.MT * Cout;
    p.tmem_cols = cols <= 32 ? 32 : cols <= 64 ? 64 : cols <= 128 ? 128 : cols <= 256 ? 256 : 512;

    static PerDeviceFlag attr_done_pd; bool& attr_done = attr_done_pd.cur();
    if (!attr_done) {
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(tc::wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 201 * 1024));
        attr_done = true;
    }
    dim3 grid(pl.groups, pl.splits);
    tc::wgrad_tc_kernel<<<grid, 192, pl.smem_bytes, s>>>(tX1, tX2, tDY, p);
    MRISR_LAUNCH_CHECK();
    const int64_t total = (int64_t)p.Mtot * Cout;
    if (pl.splits >= 16)
        tc::wgrad_tc_reduce_kernel<<<(unsigned)ceil_div64(total, 32), 256, 0, s>>>(a.partial, a.dw, p.Mtot, Cout, Cin, p.taps, pl.splits);
    else
        tc::wgrad_tc_reduce_serial_kernel<<<(unsigned)ceil_div64(total, 256), 256, 0, s>>>(a.partial, a.dw, p.Mtot, Cout, Cin, p.taps, pl.splits);
    MRISR_LAUNCH_CHECK();
    return 0;
}

}  // namespace mrisr
