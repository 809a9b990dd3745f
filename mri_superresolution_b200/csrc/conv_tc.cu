// tcgen05 implicit-GEMM convolution (3x3 pad 1 / 1x1, stride 1) over padded NHWC bf16 activations, fp32 accumulate in TMEM.
// Replaces the cuDNN convolutions behind nn.Conv2d in models/unet_model.py:29,34,72,101,152,168 for the forward pass and,
// with tap-flipped transposed weights, their data gradients (autograd; the reference has no hand-written backward).
//
// Formulation.  A padded activation [N][H+2][W+2][C] is a 2-D matrix [rows = N*(H+2)*(W+2)][C]; with a zero border a 3x3
// tap (ty, tx) is the SAME matrix shifted by (ty-1)*(W+2) + (tx-1) rows.  One CTA tile is 128 consecutive rows (pixels) x
// all Cout channels:  D[128 x Cout] = sum_{tap, k-chunk} A_shift(tap)[128 x BK] * Wt[tap][Cout x BK]^T.
// A and B tiles are fetched by TMA straight into the 32/64/128-byte swizzled K-major layout tcgen05.mma reads
// (out-of-range rows are zero-filled by TMA), so the skip-concatenation is just a second tensor map over the K loop.
// Rows that are border pixels compute garbage and are simply not stored (the zero border is an invariant).
//
// Warp roles (320 threads, 1 CTA/SM, persistent over tiles): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer,
// warps 2..9 = epilogue (TMEM -> registers -> +bias -> bf16 -> global, GroupNorm partial sums).  Two TMEM accumulator
// buffers let the epilogue of tile i overlap the MMAs of tile i+1.  Each TMEM lane quadrant is served by TWO epilogue
// warps that take alternate 16-column chunks: with one warp per scheduler the epilogue ran at the dependent-issue latency
// (~2.2 us per 128 x 64 tile measured on the 64-channel 512^2 layers, whatever the K depth) and set the pace of every
// layer with few output channels.
#include <mutex>
#include <stdlib.h>
#include "kernels.cuh"
#include "tc_common.cuh"

namespace mrisr {
namespace tc {

// ---------------------------------------------------------------- host: tensor map encoder
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    });
    return fn;
}

int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows, uint32_t box_cols,
                      uint64_t row_stride_elems) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) { set_error("cuTensorMapEncodeTiled entry point not available"); return -6; }
    const uint32_t bytes = box_cols * 2;
    CUtensorMapSwizzle sw = bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
    if (sw == CU_TENSOR_MAP_SWIZZLE_NONE || box_rows > 256 || (cols * 2) % 16 != 0 || (row_stride_elems * 2) % 16 != 0 ||
        ((uintptr_t)base & 15) != 0) {
        set_error("make_tmap_2d_bf16: unsupported box %ux%u / cols %llu / base %p", box_rows, box_cols, (unsigned long long)cols, base);
        return -6;
    }
    const cuuint64_t gdim[2] = {cols, rows};
    const cuuint64_t gstride[1] = {(row_stride_elems ? row_stride_elems : cols) * 2};     // a channel slice of a wider tensor
    const cuuint32_t box[2] = {box_cols, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return -6; }
    return 0;
}

// ---------------------------------------------------------------- device
__device__ __forceinline__ bool elect_one_c() {
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
__device__ __forceinline__ void umma_lohi_c(uint32_t tmem_d, uint32_t alo, uint32_t blo, uint32_t hi, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b64 da, db;\n"
        "setp.ne.b32 p, %5, 0;\n"
        "mov.b64 da, {%1, %3};\n"
        "mov.b64 db, {%2, %3};\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n"
        "}\n"
        ::"r"(tmem_d), "r"(alo), "r"(blo), "r"(hi), "r"(idesc), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ void umma_lohi2_c(uint32_t tmem_d, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi, uint32_t idesc,
                                             uint32_t accumulate) {
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

struct ConvTcParams {
    __nv_bfloat16* dst;     // padded origin of the output tensor (Cout = BN channels per row)
    const float* bias;      // [BN] or nullptr
    float* stat_part;       // [num_tiles][2][8][2] GroupNorm partial (sum, sumsq) or nullptr
    int chunks1, chunks2;   // K chunks of BK channels taken from source 1 / source 2
    int taps;               // 1 or 9
    int pw;                 // padded row pitch W + 2
    int S;                  // padded pixels per sample (H + 2) * (W + 2)
    int H, W;
    int total_rows;         // N * S
    int num_tiles;          // M tiles x n_blocks
    int n_blocks;           // Cout / BN (> 1 only for Cout = 512: two 256-column blocks per pixel tile)
    int ldc;                // channels per output row (= Cout)
    int relu;               // epilogue: out = max(out, 0)                         (VGG conv + ReLU, utils/losses.py:95)
    const __nv_bfloat16* mask;   // epilogue: out = mask > 0 ? out : 0, same layout as dst (ReLU backward fused into dgrad)
    int stages;             // WRES kernels: pipeline depth that fits beside the resident weights (<= Cfg::STAGES)
};

// MERGE3: one A tile of 128 + 2 pixel rows per (vertical tap, K chunk) serves the three horizontal taps - the operand
// descriptor start simply moves by one 128-byte row per tap - so the activation tile is fetched from L2 3x instead of 9x
// (the 64-channel layers were L2-bound at 9x: 64>64@128^2 0.19 -> 0.107 ms).  The 128-byte swizzle is a function of the
// absolute shared-memory address, so a row-shifted start needs NO base-offset in the descriptor (measured on B200: base
// offset 0 is bit-correct, base offset = row phase is wrong).  BK = 64 only.
template <int BN, int BK, bool MERGE3>
struct ConvTcCfg {
    static constexpr int BM = 128;
    static constexpr int A_ROWS = MERGE3 ? BM + 2 : BM;
    static constexpr int A_BYTES = ((A_ROWS * BK * 2) + 1023) / 1024 * 1024;
    static constexpr int A_TX = A_ROWS * BK * 2;                 // bytes one A load delivers
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int NB = MERGE3 ? 3 : 1;                    // weight tiles per stage
    static constexpr int STAGE_BYTES = A_BYTES + NB * B_BYTES;
    static constexpr int STAGES_RAW = (200 * 1024) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024;
    static constexpr int SWZ = BK * 2;
    static constexpr uint32_t TMEM_COLS = 2 * BN <= 32 ? 32 : 2 * BN <= 64 ? 64 : 2 * BN <= 128 ? 128 : 2 * BN <= 256 ? 256 : 512;
};

// WRES (MERGE3 only): all 9 * K-chunks weight tiles of the layer are fetched ONCE per CTA and stay in shared memory; a pipeline
// stage then carries only the activation tile.  For the 64-channel layers the per-tile weight reload was 60% of the
// L2 -> shared-memory traffic (3 x 24.6 KB of 3 x 41 KB per 128-pixel tile) and that traffic, not the MMA, set the pace
// (VGG conv1_2 64>64@512^2: 543 TFLOP/s).  Usable while 9 * Cin * Cout * 2 B leaves room for >= 3 stages.
constexpr int kConvTcThreads = 320;

template <int BN, int BK, bool MERGE3, bool WRES = false>
__global__ void __launch_bounds__(kConvTcThreads, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmA2,
               const __grid_constant__ CUtensorMap tmB, const ConvTcParams p) {
    static_assert(!WRES || MERGE3, "resident weights are implemented for the MERGE3 tile loop");
    using Cfg = ConvTcCfg<BN, BK, MERGE3>;
    constexpr int STAGES = WRES ? 8 : Cfg::STAGES;      // WRES: up to 8, the launcher passes what fits (p.stages)
    constexpr int STAGE_BYTES = WRES ? Cfg::A_BYTES : Cfg::STAGE_BYTES;
    constexpr uint32_t LT = layout_type_for(Cfg::SWZ);
    constexpr uint32_t SBO = 8 * Cfg::SWZ;
    constexpr uint32_t IDESC = make_idesc_bf16(128, BN, 0, 0);

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[STAGES];
    __shared__ __align__(8) uint64_t empty_bar[STAGES];
    __shared__ __align__(8) uint64_t tfull_bar[2];
    __shared__ __align__(8) uint64_t tempty_bar[2];
    __shared__ __align__(8) uint64_t wfull_bar;
    __shared__ uint32_t s_tmem;
    __shared__ float s_part[2][8][2][16];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nstages = WRES ? p.stages : STAGES;
    const int cpt = p.chunks1 + p.chunks2;            // K chunks per tap
    uint8_t* const wres = smem;                       // WRES: [tap][chunk] weight tiles, then the activation stages
    if constexpr (WRES) smem += 9 * cpt * Cfg::B_BYTES;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmA1);
        tma_prefetch_desc(&tmA2);
        tma_prefetch_desc(&tmB);
        mbar_init(&wfull_bar, 1);
        for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 256); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(&s_tmem, Cfg::TMEM_COLS);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = s_tmem;
    const int num_kb = (MERGE3 ? 3 : p.taps) * cpt;   // pipeline stages per tile

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            if constexpr (WRES) {
                mbar_arrive_expect_tx(&wfull_bar, (uint32_t)(9 * cpt * Cfg::B_BYTES));
                for (int tap = 0; tap < 9; ++tap)
                    for (int c = 0; c < cpt; ++c)
                        tma_load_2d(wres + (tap * cpt + c) * Cfg::B_BYTES, &tmB, &wfull_bar, c * BK, tap * p.ldc);
            }
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
                const int m0 = (tile / p.n_blocks) * 128;
                const int nb = tile % p.n_blocks;
                if constexpr (MERGE3) {
                    for (int dy = 0; dy < 3; ++dy) {
                        const int row0 = m0 + (dy - 1) * p.pw - 1;             // first of the 130 rows: tap (dy, dx = -1)
                        for (int c = 0; c < cpt; ++c) {
                            mbar_wait(&empty_bar[stage], phase ^ 1);
                            mbar_arrive_expect_tx(&full_bar[stage], WRES ? Cfg::A_TX : Cfg::A_TX + 3 * Cfg::B_BYTES);
                            uint8_t* a_dst = smem + stage * STAGE_BYTES;
                            if (c < p.chunks1) tma_load_2d(a_dst, &tmA1, &full_bar[stage], c * BK, row0);
                            else tma_load_2d(a_dst, &tmA2, &full_bar[stage], (c - p.chunks1) * BK, row0);
                            if constexpr (!WRES) {
#pragma unroll
                                for (int dx = 0; dx < 3; ++dx)
                                    tma_load_2d(a_dst + Cfg::A_BYTES + dx * Cfg::B_BYTES, &tmB, &full_bar[stage], c * BK,
                                                (dy * 3 + dx) * p.ldc + nb * BN);
                            }
                            if (++stage == nstages) { stage = 0; phase ^= 1; }
                        }
                    }
                } else {
                    for (int tap = 0; tap < p.taps; ++tap) {
                        const int shift = p.taps == 9 ? (tap / 3 - 1) * p.pw + (tap % 3 - 1) : 0;
                        for (int c = 0; c < cpt; ++c) {
                            mbar_wait(&empty_bar[stage], phase ^ 1);
                            mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
                            uint8_t* a_dst = smem + stage * Cfg::STAGE_BYTES;
                            if (c < p.chunks1) tma_load_2d(a_dst, &tmA1, &full_bar[stage], c * BK, m0 + shift);
                            else tma_load_2d(a_dst, &tmA2, &full_bar[stage], (c - p.chunks1) * BK, m0 + shift);
                            tma_load_2d(a_dst + Cfg::A_BYTES, &tmB, &full_bar[stage], c * BK, tap * p.ldc + nb * BN);
                            if (++stage == nstages) { stage = 0; phase ^= 1; }
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: warp-uniform loop, one elected lane issues; descriptor = constant hi word + (base lo + one add) =====
        constexpr uint32_t DESC_HI = ((SBO >> 4) & 0x3FFFu) | (1u << 14) | ((LT & 7u) << 29);
        const uint32_t a_lo0 = (1u << 16) | (smem_u32(smem) >> 4);                       // LBO = 16 B (unused for swizzled K-major)
        const uint32_t b_lo0 = WRES ? (1u << 16) | (smem_u32(wres) >> 4) : a_lo0 + (Cfg::A_BYTES >> 4);
        int stage = 0;
        uint32_t phase = 0;
        int it = 0;
        if constexpr (WRES) mbar_wait(&wfull_bar, 0);
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
            const int buf = it & 1;
            mbar_wait(&tempty_bar[buf], ((it >> 1) & 1) ^ 1);
            tc_fence_after_sync();
            const uint32_t d = tmem_base + (uint32_t)(buf * BN);
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after_sync();
                if (elect_one_c()) {
                    const uint32_t so = (uint32_t)stage * (uint32_t)(STAGE_BYTES >> 4);
                    if constexpr (MERGE3) {
                        // weight tile of (dy, dx, chunk): in the stage (streamed) or at [tap][chunk] of the resident block
                        uint32_t bo = so;
                        if constexpr (WRES) {
                            const int dy = kb / cpt, c = kb - dy * cpt;
                            bo = (uint32_t)((dy * 3 * cpt + c) * (Cfg::B_BYTES >> 4));
                        }
                        const uint32_t bdx = WRES ? (uint32_t)(cpt * (Cfg::B_BYTES >> 4)) : (uint32_t)(Cfg::B_BYTES >> 4);
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx)
#pragma unroll
                            for (int k = 0; k < BK / 16; ++k)       // A start moves one 128-byte row per dx (base-offset field stays 0)
                                umma_lohi2_c(d, a_lo0 + so + (uint32_t)(dx * (BK * 2 / 16) + k * 2), DESC_HI,
                                             b_lo0 + bo + (uint32_t)dx * bdx + (uint32_t)(k * 2), DESC_HI, IDESC,
                                             (kb | dx | k) != 0 ? 1u : 0u);
                    } else {
#pragma unroll
                        for (int k = 0; k < BK / 16; ++k)
                            umma_lohi_c(d, a_lo0 + so + (uint32_t)(k * 2), b_lo0 + so + (uint32_t)(k * 2), DESC_HI, IDESC, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(&empty_bar[stage]);          // frees the smem slot once these MMAs have read it
                }
                if (++stage == nstages) { stage = 0; phase ^= 1; }
            }
            if (elect_one_c()) umma_commit(&tfull_bar[buf]);    // accumulator complete -> epilogue
        }
    } else {
        // ===== epilogue: 8 warps; warp w may only touch TMEM lanes [32*(w%4), 32*(w%4)+32); half = which 16-column chunks =====
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        constexpr int GS = BN / 8;                           // channels per GroupNorm group
        constexpr int NCH = BN / 32 > 0 ? BN / 32 : 1;       // 16-column chunks per warp
        const bool do_stats = p.stat_part != nullptr;
        const int ph = p.H + 2;
        // (n, y, x) of this thread's row, advanced incrementally from tile to tile (one division pair per CTA, not per tile)
        int rn, ry, rx, dn, dy_, dx_;
        {
            const int row0 = (blockIdx.x / p.n_blocks) * 128 + q * 32 + lane;
            rn = row0 / p.S;
            const int rem = row0 - rn * p.S;
            ry = rem / p.pw; rx = rem - ry * p.pw;
            const int step = (int)gridDim.x * 128;           // rows between this CTA's consecutive tiles (n_blocks == 1)
            dn = step / p.S;
            const int srem = step - dn * p.S;
            dy_ = srem / p.pw; dx_ = srem - dy_ * p.pw;
        }
        int it = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
            const int buf = it & 1;
            const int mtile = tile / p.n_blocks, nb = tile - mtile * p.n_blocks;
            const int row = mtile * 128 + q * 32 + lane;
            int n, y, x;
            if (p.n_blocks == 1) {
                n = rn; y = ry; x = rx;
                rx += dx_; if (rx >= p.pw) { rx -= p.pw; ++ry; }
                ry += dy_; if (ry >= ph) { ry -= ph; ++rn; }
                rn += dn;
            } else {
                n = row / p.S;
                const int rem = row - n * p.S;
                y = rem / p.pw; x = rem - y * p.pw;
            }
            const bool valid = row < p.total_rows && y >= 1 && y <= p.H && x >= 1 && x <= p.W;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BN);
            __nv_bfloat16* out = p.dst + (int64_t)row * p.ldc + nb * BN;
            const __nv_bfloat16* msk = p.mask ? p.mask + (int64_t)row * p.ldc + nb * BN : nullptr;
            const float* bias = p.bias ? p.bias + nb * BN : nullptr;
            float s1[8], s2[8];
#pragma unroll
            for (int g = 0; g < 8; ++g) s1[g] = s2[g] = 0.f;
            mbar_wait(&tfull_bar[buf], (it >> 1) & 1);
            tc_fence_after_sync();
#pragma unroll
            for (int j0 = 0; j0 < NCH; j0 += 2) {
                // this warp's chunks are c0 = (2 * j + half) * 16; two TMEM loads in flight per wait
                uint32_t raw[2][16];
#pragma unroll
                for (int jj = 0; jj < 2; ++jj) {
                    const int c0 = (2 * (j0 + jj) + half) * 16;
                    if (j0 + jj < NCH && c0 < BN) tmem_ld16_nowait(taddr + c0, raw[jj]);
                }
                tmem_ld_wait();
#pragma unroll
                for (int jj = 0; jj < 2; ++jj) {
                    const int c0 = (2 * (j0 + jj) + half) * 16;
                    if (!(j0 + jj < NCH && c0 < BN)) continue;
                    float v[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(raw[jj][i]);
                    if (bias) {
                        // the bias is a slice of the flat parameter arena: 16-byte aligned for the VGG stack, at an odd
                        // element offset in the U-Net arena (alpha comes first) -> vector loads only when aligned
                        if ((reinterpret_cast<uintptr_t>(bias) & 15) == 0) {
#pragma unroll
                            for (int i = 0; i < 16; i += 4) {
                                const float4 b4 = __ldg(reinterpret_cast<const float4*>(bias + c0 + i));
                                v[i] += b4.x; v[i + 1] += b4.y; v[i + 2] += b4.z; v[i + 3] += b4.w;
                            }
                        } else {
#pragma unroll
                            for (int i = 0; i < 16; ++i) v[i] += __ldg(bias + c0 + i);
                        }
                    }
                    if (p.relu) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
                    }
                    if (msk && valid) {
                        float m0[8], m1[8];
                        load8(msk + c0, m0);
                        load8(msk + c0 + 8, m1);
#pragma unroll
                        for (int i = 0; i < 8; ++i) { v[i] = m0[i] > 0.f ? v[i] : 0.f; v[8 + i] = m1[i] > 0.f ? v[8 + i] : 0.f; }
                    }
                    if (valid) {
                        if (do_stats) {
                            // group of column c0 + i: (c0 + i) / GS with c0 = 32 * j + 16 * half; GS >= 2
#pragma unroll
                            for (int i = 0; i < 16; ++i) {
                                const int GL = (32 * (j0 + jj) + i) / GS, GH0 = (32 * (j0 + jj) + 16 + i) / GS;   // half 0 / half 1
                                const int gl = GL < 8 ? GL : 7, gh = GH0 < 8 ? GH0 : 7;
                                if (gl == gh) {
                                    s1[gl] += v[i];
                                    s2[gl] = fmaf(v[i], v[i], s2[gl]);
                                } else if (half == 0) {
                                    s1[gl] += v[i];
                                    s2[gl] = fmaf(v[i], v[i], s2[gl]);
                                } else {
                                    s1[gh] += v[i];
                                    s2[gh] = fmaf(v[i], v[i], s2[gh]);
                                }
                            }
                        }
                        // one 32-byte store per thread: two 16-byte stores to the same sector reach L2 as two partial-sector
                        // writes (rows are >= 32 B apart, so nothing else merges them), which bounded the Cout <= 64 layers
                        st_global_256(out + c0, pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                      pack_bf16x2(v[6], v[7]), pack_bf16x2(v[8], v[9]), pack_bf16x2(v[10], v[11]),
                                      pack_bf16x2(v[12], v[13]), pack_bf16x2(v[14], v[15]));
                    }
                }
            }
            tc_fence_before_sync();
            mbar_arrive(&tempty_bar[buf]);                   // TMEM buffer may be overwritten by tile it + 2
            if (p.stat_part) {
                const int n0 = (tile * 128) / p.S;
                const int slot = n - n0;                     // 0 or 1 (S >= 128)
                const int wslot = half * 4 + q;
#pragma unroll
                for (int sl = 0; sl < 2; ++sl) {
                    const bool mine = valid && slot == sl;
                    if (__ballot_sync(0xffffffffu, mine) != 0u) {
                        float w[16];
#pragma unroll
                        for (int g = 0; g < 8; ++g) { w[2 * g] = mine ? s1[g] : 0.f; w[2 * g + 1] = mine ? s2[g] : 0.f; }
                        warp_reduce16(w, lane);
                        if ((lane & 1) == 0) s_part[it & 1][wslot][sl][lane >> 1] = w[0];
                    } else if (lane < 16) {
                        s_part[it & 1][wslot][sl][lane] = 0.f;
                    }
                }
                asm volatile("bar.sync 1, 256;" ::: "memory");
                if (warp == 2) {
                    const int sl = lane >> 4, idx = lane & 15;
                    float t = 0.f;
#pragma unroll
                    for (int w8 = 0; w8 < 8; ++w8) t += s_part[it & 1][w8][sl][idx];
                    p.stat_part[((int64_t)tile * 2 + sl) * 16 + idx] = t;
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
    }
}

// stats[n][g] = (mean, rstd) from the per-tile partials of conv_tc_kernel
__global__ void __launch_bounds__(256) gn_finalize_tc_kernel(const float* __restrict__ part, float* __restrict__ stats, int S,
                                                             int num_tiles, double count) {
    __shared__ double s_acc[16][16];
    const int n = blockIdx.x;
    const int idx = threadIdx.x & 15, tl = threadIdx.x >> 4;
    const int t_first = (int)(((int64_t)n * S) / 128), t_last = (int)((((int64_t)n + 1) * S - 1) / 128);
    double acc = 0.0;
    for (int t = t_first + tl; t <= t_last && t < num_tiles; t += 16) {
        const int sl = n - (int)(((int64_t)t * 128) / S);
        if (sl == 0 || sl == 1) acc += (double)part[((int64_t)t * 2 + sl) * 16 + idx];
    }
    s_acc[tl][idx] = acc;
    __syncthreads();
    if (threadIdx.x < kGroups) {
        const int g = threadIdx.x;
        double a = 0.0, b = 0.0;
        for (int j = 0; j < 16; ++j) { a += s_acc[j][2 * g]; b += s_acc[j][2 * g + 1]; }
        const double mean = a / count;
        double var = b / count - mean * mean;
        if (var < 0.0) var = 0.0;
        stats[(n * kGroups + g) * 2] = (float)mean;
        stats[(n * kGroups + g) * 2 + 1] = (float)(1.0 / sqrt(var + (double)kGnEps));
    }
}

// fp32 NCHW [Cout][Cin][ks][ks] -> wt_fwd[tap][Cout][Cin] and wt_dgrad[taps-1-tap][Cin][Cout] (both K-major bf16)
__global__ void pack_weights_tc_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wf,
                                       __nv_bfloat16* __restrict__ wd, int Cout, int Cin, int taps) {
    const int64_t total = (int64_t)Cout * Cin * taps;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int tap = (int)(i % taps);
        const int ci = (int)((i / taps) % Cin);
        const int co = (int)(i / ((int64_t)taps * Cin));
        const __nv_bfloat16 v = __float2bfloat16_rn(w[i]);
        wf[((int64_t)tap * Cout + co) * Cin + ci] = v;
        if (wd) wd[((int64_t)(taps - 1 - tap) * Cin + ci) * Cout + co] = v;
    }
}

constexpr int kMaxSmem = 222 * 1024;        // dynamic part; the static barriers and partial-sum scratch take 3.2 KB of the 227 KB

// resident-weight variant: shared memory = weights + p.stages activation tiles
template <int BN, int BK>
static int launch_wres(const CUtensorMap& tA1, const CUtensorMap& tA2, const CUtensorMap& tB, const ConvTcParams& p, int grid,
                       cudaStream_t s) {
    using Cfg = ConvTcCfg<BN, BK, true>;
    static PerDeviceFlag attr_done_pd; bool& attr_done = attr_done_pd.cur();
    if (!attr_done) {
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(conv_tc_kernel<BN, BK, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem));
        attr_done = true;
    }
    const int smem = 9 * (p.chunks1 + p.chunks2) * Cfg::B_BYTES + p.stages * Cfg::A_BYTES + 1024;
    conv_tc_kernel<BN, BK, true, true><<<grid, kConvTcThreads, smem, s>>>(tA1, tA2, tB, p);
    MRISR_LAUNCH_CHECK();
    return 0;
}

// pipeline depth of the resident-weight variant for this layer, 0 = does not fit / not worth it
template <int BN, int BK>
static int wres_stages(const ConvTcParams& p, int grid) {
    using Cfg = ConvTcCfg<BN, BK, true>;
    if (p.n_blocks != 1 || p.num_tiles < 4 * grid) return 0;      // the one-off weight fetch must amortise
    // measured (profiles/r01_conv_tc_epilogue.md): a gain where the weight tiles dominate a stage (Cout = 128: 3 x 16 KB beside
    // a 17 KB activation tile; 32-channel chunks), a loss for Cout <= 64 with 64-channel chunks (64>64, 128>64, 64>32)
    if (!(BN >= 128 || BK == 32)) return 0;
    const int left = kMaxSmem - 1024 - 9 * (p.chunks1 + p.chunks2) * Cfg::B_BYTES;
    const int st = left / Cfg::A_BYTES;
    return st < 3 ? 0 : (st > 8 ? 8 : st);
}

template <int BN, int BK, bool MERGE3>
static int launch_cfg(const CUtensorMap& tA1, const CUtensorMap& tA2, const CUtensorMap& tB, const ConvTcParams& p, int grid,
                      cudaStream_t s) {
    using Cfg = ConvTcCfg<BN, BK, MERGE3>;
    if constexpr (MERGE3) {
        const int st = wres_stages<BN, BK>(p, grid);
        if (st > 0) {
            ConvTcParams q = p;
            q.stages = st;
            return launch_wres<BN, BK>(tA1, tA2, tB, q, grid, s);
        }
    }
    static PerDeviceFlag attr_done_pd; bool& attr_done = attr_done_pd.cur();
    if (!attr_done) {
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(conv_tc_kernel<BN, BK, MERGE3>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        attr_done = true;
    }
    conv_tc_kernel<BN, BK, MERGE3><<<grid, kConvTcThreads, Cfg::SMEM_BYTES, s>>>(tA1, tA2, tB, p);
    MRISR_LAUNCH_CHECK();
    return 0;
}

template <int BN>
static int launch_bn(int bk, bool merge3, const CUtensorMap& tA1, const CUtensorMap& tA2, const CUtensorMap& tB, const ConvTcParams& p,
                     int grid, cudaStream_t s) {
    if constexpr (BN <= 128) {
        if (merge3) return bk == 64 ? launch_cfg<BN, 64, true>(tA1, tA2, tB, p, grid, s)
                         : bk == 32 ? launch_cfg<BN, 32, true>(tA1, tA2, tB, p, grid, s)
                                    : launch_cfg<BN, 16, true>(tA1, tA2, tB, p, grid, s);
    }
    switch (bk) {
        case 64: return launch_cfg<BN, 64, false>(tA1, tA2, tB, p, grid, s);
        case 32: return launch_cfg<BN, 32, false>(tA1, tA2, tB, p, grid, s);
        default: return launch_cfg<BN, 16, false>(tA1, tA2, tB, p, grid, s);
    }
}

int num_sms() {
    static int cache[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    int& n = cache[dev];                       // per device: a process may drive GPUs with different SM counts
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    }
    return n;
}

}  // namespace tc

using bf16 = __nv_bfloat16;

// MRISR_CONV_TC_MERGE3=0 disables the 3-taps-per-tile path (debugging / A-B timing)
static bool conv_tc_merge3_enabled() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("MRISR_CONV_TC_MERGE3"); v = (e && e[0] == '0') ? 0 : 1; }
    return v != 0;
}

bool conv_tc_supported(int H, int W, int C1, int C2, int Cout, int ks) {
    const bool cout_ok = Cout == 16 || Cout == 32 || Cout == 64 || Cout == 128 || Cout == 256 || Cout == 512;
    // (H+2)*(W+2) >= 128 is only needed by the GroupNorm-statistics epilogue (at most two samples per tile); the launcher
    // rejects stat_part on smaller images
    return cout_ok && C1 >= 16 && C1 % 16 == 0 && C2 % 16 == 0 && (ks == 1 || ks == 3) && H >= 1 && W >= 1;
}

int conv_tc_stat_elems(int N, int H, int W) {
    const int64_t rows = (int64_t)N * (H + 2) * (W + 2);
    return (int)(((rows + 127) / 128) * 32);
}

int launch_conv_tc(const ConvArgs<bf16>& a, cudaStream_t s) {
    const int C1 = a.src1.C, C2 = a.src2.p ? a.src2.C : 0, Cin = C1 + C2, Cout = a.dst.C;
    const int H = a.dst.H, W = a.dst.W, N = a.dst.N;
    if (!conv_tc_supported(H, W, C1, C2, Cout, a.ks)) {
        set_error("conv_tc: unsupported shape C1=%d C2=%d Cout=%d ks=%d %dx%d", C1, C2, Cout, a.ks, H, W);
        return -2;
    }
    const char* tag = "conv_tc";
    char tbuf[96];
    if (prof_on()) {
        snprintf(tbuf, sizeof(tbuf), "conv_tc:k%d:%d>%d@%dx%d", a.ks, Cin, Cout, H, W);
        tag = prof_intern(tbuf);
    }
    ProfScope prof_(tag, s);
    int bk = 64;
    if (C1 % 64 || C2 % 64) bk = (C1 % 32 || C2 % 32) ? 16 : 32;
    const int64_t rows = (int64_t)N * (H + 2) * (W + 2);
    const int taps = a.ks * a.ks;
    const int BNsel = Cout > 256 ? 256 : Cout;
    // 32- / 16-channel chunks use the 64- / 32-byte swizzle: the same row-shifted descriptor start works
    const bool merge3 = conv_tc_merge3_enabled() && taps == 9 && BNsel <= 128;
    const uint32_t a_rows = merge3 ? 130u : 128u;
    CUtensorMap tA1, tA2, tB;
    int rc = tc::make_tmap_2d_bf16(&tA1, a.src1.p, (uint64_t)rows, (uint64_t)C1, a_rows, (uint32_t)bk);
    if (rc) return rc;
    if (C2) { rc = tc::make_tmap_2d_bf16(&tA2, a.src2.p, (uint64_t)rows, (uint64_t)C2, a_rows, (uint32_t)bk); if (rc) return rc; }
    else tA2 = tA1;
    const int BN = Cout > 256 ? 256 : Cout;
    rc = tc::make_tmap_2d_bf16(&tB, a.w, (uint64_t)taps * Cout, (uint64_t)Cin, (uint32_t)BN, (uint32_t)bk);
    if (rc) return rc;
    if (a.stat_part && ((int64_t)(H + 2) * (W + 2) < 128 || Cout > 256)) {
        set_error("conv_tc: GroupNorm statistics need (H+2)*(W+2) >= 128 and Cout <= 256");
        return -2;
    }
    tc::ConvTcParams p;
    p.dst = a.dst.p; p.bias = a.bias; p.stat_part = a.stat_part;
    p.n_blocks = Cout / BN; p.ldc = Cout; p.relu = a.relu ? 1 : 0; p.mask = a.mask; p.stages = 0;
    p.chunks1 = C1 / bk; p.chunks2 = C2 / bk; p.taps = taps;
    p.pw = W + 2; p.S = (H + 2) * (W + 2); p.H = H; p.W = W;
    p.total_rows = (int)rows; p.num_tiles = (int)((rows + 127) / 128) * p.n_blocks;
    const int grid = p.num_tiles < tc::num_sms() ? p.num_tiles : tc::num_sms();
    switch (BN) {
        case 16: return tc::launch_bn<16>(bk, merge3, tA1, tA2, tB, p, grid, s);
        case 32: return tc::launch_bn<32>(bk, merge3, tA1, tA2, tB, p, grid, s);
        case 64: return tc::launch_bn<64>(bk, merge3, tA1, tA2, tB, p, grid, s);
        case 128: return tc::launch_bn<128>(bk, merge3, tA1, tA2, tB, p, grid, s);
        default: return tc::launch_bn<256>(bk, merge3, tA1, tA2, tB, p, grid, s);
    }
}

int launch_gn_finalize_tc(const float* part, float* stats, int N, int H, int W, double count, cudaStream_t s) {
    ProfScope prof_("gn_finalize", s);
    const int S = (H + 2) * (W + 2);
    const int num_tiles = (int)(((int64_t)N * S + 127) / 128);
    tc::gn_finalize_tc_kernel<<<N, 256, 0, s>>>(part, stats, S, num_tiles, count);
    MRISR_LAUNCH_CHECK();
    return 0;
}

int launch_pack_weights_tc(const float* w, bf16* wf, bf16* wd, int Cout, int Cin, int ks, cudaStream_t s) {
    ProfScope prof_("pack_weights", s);
    const int64_t total = (int64_t)Cout * Cin * ks * ks;
    int blocks = (int)ceil_div64(total, 256);
    if (blocks > 1184) blocks = 1184;
    tc::pack_weights_tc_kernel<<<blocks, 256, 0, s>>>(w, wf, wd, Cout, Cin, ks * ks);
    MRISR_LAUNCH_CHECK();
    return 0;
}

}  // namespace mrisr
