// tcgen05 weight gradient over padded NHWC bf16 activations with the three horizontal taps served by ONE fetched tile
// (second generation of wgrad_tc.cu for inputs chunked by 64 or 32 channels, Cout = 64 / 128):
//   dW[tap][ci][co] = sum_rows X[row + shift(tap)][ci] * dY[row][co]        (autograd of nn.Conv2d, models/unet_model.py:29..172)
// wgrad_tc.cu fetches one [64 pixels x 64 channels] box of X per (tap, chunk) and K block: X crosses L2 -> shared memory 9x
// and the kernel is bound by that traffic (ncu: 0.8 TB/s DRAM, profiles/r01_launches_step.md).  Here a box of 64 + 2 pixel
// rows is fetched once per (vertical tap, chunk); the operand descriptor start moves by one 128-byte row per horizontal tap
// (the 128-byte swizzle is a function of the absolute shared-memory address, so no re-layout is needed - measured in
// conv_tc.cu MERGE3), i.e. 3x instead of 9x.  A CTA group owns SG consecutive slabs (dy, chunk) and 3 * ceil(SG / 2)
// accumulators [128 x Cout] in TMEM; CTAs of a group split the pixel range; wgrad_tc3_reduce_kernel sums the splits in a
// fixed order into the reference's [Cout][Cin][kh][kw] fp32 layout.
#include "kernels.cuh"
#include "tc_common.cuh"

namespace mrisr {
namespace tc {

int num_sms();

__device__ __forceinline__ bool elect_one_3() {
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
__device__ __forceinline__ void umma_lohi_3(uint32_t tmem_d, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi,
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

constexpr int kBKP3 = 64;                       // pixels (GEMM K) per pipeline stage
constexpr int kMaxStages3 = 8;
// a slab = (64 + 2) pixel rows x CS channels (CS = 64: 128-byte rows / 128B swizzle; CS = 32: 64-byte rows / 64B swizzle),
// padded to a multiple of 1024 bytes; an M = 128 tile is 128 / CS consecutive slabs (LBO = slab stride)
__host__ __device__ constexpr uint32_t slab_bytes3(int cs) { return ((66u * (uint32_t)cs * 2u) + 1023u) / 1024u * 1024u; }

struct Wgrad3Params {
    float* partial;               // [splits][groups][3 * npairs][128][BN]
    int BN, SG, npairs, cpt, chunks1, nslabs, pw, CS;      // npairs = M tiles per CTA (pairs of 64-channel slabs, quads of 32)
    int num_kb, kb_per_split, stages;
    uint32_t stage_bytes, b_bytes, idesc, tmem_cols;
};

template <int CS>
__global__ void __launch_bounds__(192, 1)
wgrad_tc3_kernel(const __grid_constant__ CUtensorMap tmX1, const __grid_constant__ CUtensorMap tmX2,
                 const __grid_constant__ CUtensorMap tmDY, const Wgrad3Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[kMaxStages3];
    __shared__ __align__(8) uint64_t empty_bar[kMaxStages3];
    __shared__ __align__(8) uint64_t done_bar;
    __shared__ uint32_t s_tmem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = blockIdx.x;                                      // slab group
    const int slab0 = g * p.SG;
    const int sgv = min(p.SG, p.nslabs - slab0);                   // valid slabs of this group
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
            constexpr uint32_t slab = slab_bytes3(CS);
            const uint32_t tx_bytes = (uint32_t)sgv * (66u * (uint32_t)CS * 2u) + p.b_bytes;
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = kb_begin; kb < kb_end; ++kb) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
                uint8_t* base = smem + (size_t)stage * p.stage_bytes;
                const int row0 = kb * kBKP3;
                for (int j = 0; j < sgv; ++j) {
                    const int s = slab0 + j, dy = s / p.cpt, chunk = s - dy * p.cpt;
                    const int r = row0 + (dy - 1) * p.pw - 1;      // first of the 66 rows: tap (dy, dx = -1)
                    uint8_t* dst = base + (size_t)j * slab;
                    if (chunk < p.chunks1) tma_load_2d(dst, &tmX1, &full_bar[stage], chunk * CS, r);
                    else tma_load_2d(dst, &tmX2, &full_bar[stage], (chunk - p.chunks1) * CS, r);
                }
                uint8_t* bdst = base + (size_t)p.SG * slab;
                for (int b = 0; b < p.BN / 64; ++b)
                    tma_load_2d(bdst + (size_t)b * 8192u, &tmDY, &full_bar[stage], b * 64, row0);
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // MN-major operands.  A: CS-channel slabs, rows of CS * 2 bytes with the matching swizzle; LBO = next slab, SBO = next 8
        // pixels; a K step of 16 pixels = 2 SBO.  B (dY): 64-channel slabs with the 128-byte swizzle.
        constexpr uint32_t slab = slab_bytes3(CS);
        const uint32_t row16 = (uint32_t)CS * 2u >> 4, sbo_a = 8u * (uint32_t)CS * 2u;
        const uint32_t a_hi = (sbo_a >> 4) | (1u << 14) | (layout_type_for(CS * 2) << 29);
        const uint32_t b_hi = (1024u >> 4) | (1u << 14) | (layout_type_for(128) << 29);
        const uint32_t a_lo0 = ((slab >> 4) << 16) | (smem_u32(smem) >> 4);
        const uint32_t b_lo0 = ((8192u >> 4) << 16) | ((smem_u32(smem) + (uint32_t)p.SG * slab) >> 4);
        const uint32_t spt = 128u / (uint32_t)CS, kstep_a = 2u * sbo_a >> 4;
        const uint32_t stage16 = p.stage_bytes >> 4;
        const uint32_t bn = (uint32_t)p.BN, idesc = p.idesc;
        const int npv = (sgv + (int)spt - 1) / (int)spt;           // M tiles that hold at least one valid slab
        int stage = 0;
        uint32_t phase = 0, acc = 0;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after_sync();
            if (elect_one_3()) {
                const uint32_t a0 = a_lo0 + (uint32_t)stage * stage16, b0 = b_lo0 + (uint32_t)stage * stage16;
#pragma unroll
                for (int k = 0; k < kBKP3 / 16; ++k) {
#pragma unroll
                    for (int dx = 0; dx < 3; ++dx)
                        for (int pr = 0; pr < npv; ++pr)
                            umma_lohi_3(tmem_base + (uint32_t)(dx * p.npairs + pr) * bn,
                                        a0 + (uint32_t)pr * (spt * slab >> 4) + (uint32_t)dx * row16 + (uint32_t)k * kstep_a, a_hi,
                                        b0 + (uint32_t)(k * 128), b_hi, idesc, acc);
                    acc = 1u;
                }
                umma_commit(&empty_bar[stage]);
            }
            acc = 1u;
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        if (elect_one_3()) umma_commit(&done_bar);
    } else {
        const int q = warp & 3;
        mbar_wait(&done_bar, 0);
        tc_fence_after_sync();
        const int ntiles = 3 * p.npairs;
        const int row = q * 32 + lane;
        for (int t = 0; t < ntiles; ++t) {
            if ((t % p.npairs) * (128 / CS) >= sgv) continue;    // M tile without valid slabs: never accumulated
            float* out = p.partial + ((((int64_t)blockIdx.y * gridDim.x + g) * ntiles + t) * 128 + row) * p.BN;
            for (int c0 = 0; c0 < p.BN; c0 += 16) {
                float v[16];
                tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(t * p.BN + c0), v);
#pragma unroll
                for (int i = 0; i < 16; i += 4)
                    *reinterpret_cast<float4*>(out + c0 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
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

// dw[co][ci][tap] = sum over splits of partial[split][group][dx * npairs + pair][(j & 1) * 64 + ci % 64][co]
// with slab s = dy * cpt + ci / 64, group = s / SG, j = s % SG, pair = j / 2
__global__ void __launch_bounds__(256) wgrad_tc3_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dw, int Cin,
                                                                int Cout, int cpt, int SG, int npairs, int groups, int splits, int CS) {
    const int total = 9 * Cin * Cout;
    const int i = blockIdx.x * 32 + (threadIdx.x & 31);
    const bool valid = i < total;
    const int iv = valid ? i : 0;
    const int co = iv % Cout, ci = (iv / Cout) % Cin, tap = iv / (Cout * Cin);
    const int dy = tap / 3, dx = tap - dy * 3;
    const int spt = 128 / CS;
    const int s = dy * cpt + ci / CS, g = s / SG, j = s - g * SG;
    const int tile = dx * npairs + j / spt, row = (j % spt) * CS + ci % CS;
    const int ntiles = 3 * npairs;
    const float t = sliced_partial_sum(partial + ((((int64_t)g * ntiles + tile) * 128 + row) * Cout + co),
                                       (int64_t)groups * ntiles * 128 * Cout, splits, valid);
    if (valid && threadIdx.x < 32) dw[((int64_t)co * Cin + ci) * 9 + tap] = t;
}

struct Wgrad3Plan { int SG, npairs, groups, splits, kb_per_split, num_kb, stages, nslabs, cpt, CS; uint32_t stage_bytes, smem_bytes, tmem_cols; bool ok; };

static Wgrad3Plan plan3(int N, int H, int W, int C1, int C2, int Cout) {
    Wgrad3Plan pl;
    pl.ok = false;
    if (!(Cout == 64 || Cout == 128)) return pl;
    if (C1 % 64 == 0 && C2 % 64 == 0) pl.CS = 64;
    else if (C1 % 32 == 0 && C2 % 32 == 0) pl.CS = 32;
    else return pl;
    const int spt = 128 / pl.CS;                       // slabs per M = 128 tile
    pl.cpt = (C1 + C2) / pl.CS;
    pl.nslabs = 3 * pl.cpt;
    const int max_tiles = 512 / Cout;                  // TMEM columns
    const int np_max = max_tiles / 3;                  // M tiles per CTA (each needs 3 accumulators)
    if (np_max < 1) return pl;
    const int sg_max = spt * np_max;
    pl.groups = ceil_div(pl.nslabs, sg_max);
    pl.SG = ceil_div(pl.nslabs, pl.groups);
    pl.npairs = ceil_div(pl.SG, spt);
    const uint32_t slab = slab_bytes3(pl.CS);
    pl.stage_bytes = (uint32_t)pl.SG * slab + (uint32_t)Cout * kBKP3 * 2u;
    int st = (int)((200u * 1024u) / pl.stage_bytes);
    pl.stages = st > kMaxStages3 ? kMaxStages3 : st;
    if (pl.stages < 2) return pl;
    pl.smem_bytes = (uint32_t)pl.stages * pl.stage_bytes + 1024u + (uint32_t)(spt - 1) * slab;     // + reach of the padding slabs of a partial tile
    const int64_t rows = (int64_t)N * (H + 2) * (W + 2);
    pl.num_kb = (int)ceil_div64(rows, kBKP3);
    int sp = num_sms() / pl.groups;
    if (sp < 1) sp = 1;
    if (sp > pl.num_kb) sp = pl.num_kb;
    pl.kb_per_split = ceil_div(pl.num_kb, sp);
    pl.splits = ceil_div(pl.num_kb, pl.kb_per_split);
    const int cols = 3 * pl.npairs * Cout;
    pl.tmem_cols = cols <= 32 ? 32 : cols <= 64 ? 64 : cols <= 128 ? 128 : cols <= 256 ? 256 : 512;
    pl.ok = cols <= 512;
    return pl;
}

}  // namespace tc

using bf16 = __nv_bfloat16;

bool wgrad_tc3_supported(int N, int H, int W, int C1, int C2, int Cout, int ks) {
    return ks == 3 && C1 >= 32 && tc::plan3(N, H, W, C1, C2, Cout).ok;
}

size_t wgrad_tc3_scratch_elems(int N, int H, int W, int C1, int C2, int Cout) {
    const tc::Wgrad3Plan pl = tc::plan3(N, H, W, C1, C2, Cout);
    if (!pl.ok) return 0;
    return (size_t)pl.splits * pl.groups * 3 * pl.npairs * 128 * Cout;
}

int launch_wgrad_tc3(const WgradArgs<bf16>& a, cudaStream_t s) {
    const int C1 = a.src1.C, C2 = a.src2.p ? a.src2.C : 0, Cin = C1 + C2, Cout = a.dy.C;
    const int H = a.dy.H, W = a.dy.W, N = a.dy.N;
    if (!wgrad_tc3_supported(N, H, W, C1, C2, Cout, a.ks)) {
        set_error("wgrad_tc3: unsupported shape C1=%d C2=%d Cout=%d ks=%d", C1, C2, Cout, a.ks);
        return -2;
    }
    const char* tag = "wgrad_tc3";
    char tbuf[96];
    if (prof_on()) {
        snprintf(tbuf, sizeof(tbuf), "wgrad_tc3:k%d:%d>%d@%dx%d", a.ks, Cin, Cout, H, W);
        tag = prof_intern(tbuf);
    }
    ProfScope prof_(tag, s);
    const tc::Wgrad3Plan pl = tc::plan3(N, H, W, C1, C2, Cout);
    const int64_t rows = (int64_t)N * (H + 2) * (W + 2);
    CUtensorMap tX1, tX2, tDY;
    int rc = tc::make_tmap_2d_bf16(&tX1, a.src1.p, (uint64_t)rows, (uint64_t)C1, tc::kBKP3 + 2, (uint32_t)pl.CS);
    if (rc) return rc;
    if (C2) { rc = tc::make_tmap_2d_bf16(&tX2, a.src2.p, (uint64_t)rows, (uint64_t)C2, tc::kBKP3 + 2, (uint32_t)pl.CS); if (rc) return rc; }
    else tX2 = tX1;
    rc = tc::make_tmap_2d_bf16(&tDY, a.dy.p, (uint64_t)rows, (uint64_t)Cout, tc::kBKP3, 64);
    if (rc) return rc;

    tc::Wgrad3Params p;
    p.partial = a.partial;
    p.BN = Cout; p.SG = pl.SG; p.npairs = pl.npairs; p.cpt = pl.cpt; p.chunks1 = C1 / pl.CS; p.nslabs = pl.nslabs; p.pw = W + 2; p.CS = pl.CS;
    p.num_kb = pl.num_kb; p.kb_per_split = pl.kb_per_split; p.stages = pl.stages;
    p.stage_bytes = pl.stage_bytes; p.b_bytes = (uint32_t)Cout * tc::kBKP3 * 2u;
    p.idesc = tc::make_idesc_bf16(128, Cout, 1, 1);
    p.tmem_cols = pl.tmem_cols;
    static PerDeviceFlag attr_done_pd; bool& attr_done = attr_done_pd.cur();
    if (!attr_done) {
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(tc::wgrad_tc3_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(tc::wgrad_tc3_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        attr_done = true;
    }
    dim3 grid(pl.groups, pl.splits);
    if (pl.CS == 64) tc::wgrad_tc3_kernel<64><<<grid, 192, pl.smem_bytes, s>>>(tX1, tX2, tDY, p);
    else tc::wgrad_tc3_kernel<32><<<grid, 192, pl.smem_bytes, s>>>(tX1, tX2, tDY, p);
    MRISR_LAUNCH_CHECK();
    const int total = 9 * Cin * Cout;
    tc::wgrad_tc3_reduce_kernel<<<ceil_div(total, 32), 256, 0, s>>>(a.partial, a.dw, Cin, Cout, pl.cpt, pl.SG, pl.npairs, pl.groups, pl.splits, pl.CS);
    MRISR_LAUNCH_CHECK();
    return 0;
}

}  // namespace mrisr
