// tcgen05 implicit-GEMM 3x3 convolution for NARROW, HIGH-RESOLUTION layers (Cin <= 64 at 256^2 / 512^2), over the
// planar-8 bf16 layout (common.cuh: C/8 planes of [pixel][8 channels]).  Same math as conv_tc.cu (forward conv of
// models/unet_model.py:29,34,101,152,168 and, with tap-flipped transposed weights, its data gradient) but organised around
// what bounds these layers: with Cout <= 64 the tensor pipe is starved by operand traffic, and re-fetching the activation
// tile once per tap (9x from L2, one TMA request per 32..128-byte row) costs ~5x the HBM time of the layer.  Here
//   * a CTA walks a contiguous run of pixel blocks (BLK = 128 or 256 pixels) and keeps a sliding window of the input in
//     shared memory: a ring of blocks per 8-channel plane, each block ONE cp.async.bulk copy, so every input byte is
//     fetched from HBM/L2 exactly once;
//   * the planar block IS the no-swizzle K-major core-matrix layout of tcgen05.mma (8 pixels x 16 B contiguous), so a tap
//     is just the same descriptor with its start address moved by shift * 16 B - no data movement per tap at all;
//     a 128-pixel mirror of the start of ring slot 0 behind the last slot keeps every 128-pixel operand window contiguous
//     across the wrap;
//   * the whole packed weight tensor (<= 72 KB) is resident in shared memory for the lifetime of the CTA;
//   * the issue loop is warp-uniform and fully unrolled over (tile, tap, K chunk): a descriptor is one integer add, so the
//     single issuing thread spends ~3 instructions per MMA (the first version spent ~220 and was issue-bound 14x over the
//     HBM time, profiles/r01_conv_ring_issue_bound.md);
//   * GroupNorm partial statistics are accumulated in registers across the CTA's whole run and flushed once per sample.
// Warp roles: warp 0 producer (bulk copies), warp 1 TMEM allocator + MMA issuer, warps 2..5 epilogue.
#include <stdlib.h>
#include "kernels.cuh"
#include "tc_common.cuh"

namespace mrisr {
namespace tc {

int num_sms();

// 1-D bulk copy global -> shared, completion counted in bytes on `bar` (addresses and size multiples of 16)
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ bool elect_one() {
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
// tcgen05.mma with descriptors given as (lo, hi) 32-bit halves: the issue loops only ever change the lo word
__device__ __forceinline__ void umma_bf16_lohi(uint32_t tmem_d, uint32_t alo, uint32_t blo, uint32_t hi, uint32_t idesc,
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

constexpr int kRingMaxSlots = 16;
constexpr int kRingTmemBufs = 4;
constexpr int kRingRecFloats = 32;      // statistics record: [0..15] (sum, sumsq) x 8 groups, [16] sample index (int bits)

struct ConvRingParams {
    const __nv_bfloat16* src1; int64_t cs1; int planes1;    // planar-8 sources (pixel 0 of plane 0, plane stride in elements)
    const __nv_bfloat16* src2; int64_t cs2;
    const __nv_bfloat16* w;         // packed [tap][Cin/8][Cout][8]
    __nv_bfloat16* dst; int64_t csd;                         // planar-8 output, Cout = BN
    const float* bias;
    float* stat_rec;                // [grid][4 warps][rec_per_warp][kRingRecFloats] or nullptr
    int rec_per_warp;
    int pw, S, H, W, total_rows, num_steps, steps_per_cta;
    int R;                          // ring slots (excluding the mirror)
    int halo_lo, halo_hi;           // blocks below / above a step's own block that its windows touch
    int sh[9];                      // halo_lo * BLK + pixel shift of tap t  (>= 0)
    uint32_t w_bytes;
};

// BN = Cout, KP = Cin / 8 (input planes), G = 128-pixel MMA tiles per step (block = 128 * G pixels)
template <int BN, int KP, int G>
__global__ void __launch_bounds__(192, 1) conv_ring_kernel(const __grid_constant__ ConvRingParams p) {
    constexpr uint32_t IDESC = make_idesc_bf16(128, BN, 0, 0);
    constexpr uint32_t TMEM_COLS = kRingTmemBufs * G * BN < 32 ? 32 : kRingTmemBufs * G * BN;
    constexpr int BLK = 128 * G;
    constexpr uint32_t BLKB = 2048u * G;                           // bytes of one block of one plane
    constexpr uint32_t DESC_HI = (128u >> 4) | (1u << 14);         // SBO = 128 B, descriptor version 1, no swizzle
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[kRingMaxSlots];
    __shared__ __align__(8) uint64_t empty_bar[kRingMaxSlots];
    __shared__ __align__(8) uint64_t tfull_bar[kRingTmemBufs];
    __shared__ __align__(8) uint64_t tempty_bar[kRingTmemBufs];
    __shared__ __align__(8) uint64_t w_bar;
    __shared__ uint32_t s_tmem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int R = p.R;
    const uint32_t PLS = (uint32_t)R * BLKB + 2048u;               // plane stride inside the ring (slots + mirror)
    uint8_t* w_smem = smem;
    uint8_t* ring = smem + ((p.w_bytes + 1023u) & ~1023u);
    const int s_begin = blockIdx.x * p.steps_per_cta;
    const int s_end = min(p.num_steps, s_begin + p.steps_per_cta);
    const int n_steps = s_end - s_begin;
    const int num_blocks = n_steps + p.halo_lo + p.halo_hi;       // blocks this CTA loads: [s_begin - halo_lo, s_end - 1 + halo_hi]

    if (threadIdx.x == 0) {
        for (int i = 0; i < R; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < kRingTmemBufs; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 4); }
        mbar_init(&w_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(&s_tmem, TMEM_COLS);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = s_tmem;

    if (warp == 0) {
        // ===== producer: weights once, then one bulk copy per (block, plane) =====
        if (lane == 0 && n_steps > 0) {
            mbar_arrive_expect_tx(&w_bar, p.w_bytes);
            for (uint32_t off = 0; off < p.w_bytes; off += 16384u) {
                const uint32_t n = min(16384u, p.w_bytes - off);
                bulk_load(w_smem + off, reinterpret_cast<const uint8_t*>(p.w) + off, n, &w_bar);
            }
            const __nv_bfloat16* g1 = p.src1 + ((int64_t)(s_begin - p.halo_lo) * BLK) * 8;
            const __nv_bfloat16* g2 = p.src2 + ((int64_t)(s_begin - p.halo_lo) * BLK) * 8;
            int slot = 0;
            uint32_t phase = 1;                                    // parity to wait for on empty_bar (first pass: free)
            for (int j = 0; j < num_blocks; ++j) {
                mbar_wait(&empty_bar[slot], phase);
                const bool mirror = slot == 0;
                mbar_arrive_expect_tx(&full_bar[slot], (uint32_t)KP * (BLKB + (mirror ? 2048u : 0u)));
                uint8_t* dst = ring + (uint32_t)slot * BLKB;
#pragma unroll
                for (int k = 0; k < KP; ++k) {
                    const __nv_bfloat16* src = k < p.planes1 ? g1 + (int64_t)k * p.cs1 : g2 + (int64_t)(k - p.planes1) * p.cs2;
                    bulk_load(dst + (uint32_t)k * PLS, src, BLKB, &full_bar[slot]);
                    if (mirror) bulk_load(dst + (uint32_t)k * PLS + (uint32_t)R * BLKB, src, 2048u, &full_bar[slot]);
                }
                g1 += BLK * 8; g2 += BLK * 8;
                if (++slot == R) { slot = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (warp-uniform loop, one elected lane issues) =====
        if (n_steps > 0) {
            mbar_wait(&w_bar, 0);
            const uint32_t ring_px = (uint32_t)R * BLK;            // ring length in pixels (= 16-byte units per plane)
            const uint32_t pls16 = PLS >> 4;
            const uint32_t a_base = (pls16 << 16) | (smem_u32(ring) >> 4);              // LBO = plane stride
            const uint32_t b_base = ((uint32_t)BN << 16) | (smem_u32(w_smem) >> 4);      // LBO = BN * 16 B (next K chunk)
            uint32_t sh[9];
#pragma unroll
            for (int t = 0; t < 9; ++t) sh[t] = (uint32_t)p.sh[t];
            const int first_need = p.halo_lo + p.halo_hi;          // highest block (relative) step 0 touches
            int ready = 0, rslot = 0;
            uint32_t rphase = 0;
            uint32_t tb = 0;                                       // (it * BLK) mod ring_px
            int eslot = 0;                                         // ring slot of block `it` (released after step it)
            for (int it = 0; it < n_steps; ++it) {
                const int need = it + first_need;
                while (ready <= need) {
                    mbar_wait(&full_bar[rslot], rphase);
                    ++ready;
                    if (++rslot == R) { rslot = 0; rphase ^= 1u; }
                }
                const int buf = it & (kRingTmemBufs - 1);
                mbar_wait(&tempty_bar[buf], ((it / kRingTmemBufs) & 1) ^ 1);
                tc_fence_after_sync();
                const bool leader = elect_one();
                if (leader) {
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        const uint32_t d = tmem_base + (uint32_t)((buf * G + g) * BN);
#pragma unroll
                        for (int tap = 0; tap < 9; ++tap) {
                            uint32_t off = tb + (uint32_t)(g * 128) + sh[tap];
                            if (off >= ring_px) off -= ring_px;
#pragma unroll
                            for (int kc = 0; kc < KP; kc += 2) {
                                const uint32_t alo = a_base + off + (uint32_t)kc * pls16;
                                const uint32_t blo = b_base + (uint32_t)((tap * KP + kc) * BN);
                                umma_bf16_lohi(d, alo, blo, DESC_HI, IDESC, (tap | kc) ? 1u : 0u);
                            }
                        }
                    }
                }
                if (leader) {
                    umma_commit(&tfull_bar[buf]);
                    umma_commit(&empty_bar[eslot]);               // block `it` is below every later window
                }
                if (++eslot == R) eslot = 0;
                tb += BLK;
                if (tb >= ring_px) tb -= ring_px;
            }
        }
    } else {
        // ===== epilogue (TMEM lane quarter = warp % 4) =====
        const int q = warp & 3;
        constexpr int GS = BN / 8;
        const int HP = p.H + 2;
        const int adv_y = 128 / p.pw, adv_x = 128 % p.pw;
        int row = s_begin * BLK + q * 32 + lane;
        int n = row / p.S;
        int y = (row - n * p.S) / p.pw;
        int x = row - n * p.S - y * p.pw;
        float s1[8], s2[8];
#pragma unroll
        for (int g8 = 0; g8 < 8; ++g8) s1[g8] = s2[g8] = 0.f;
        int cur_n = -1, nrec = 0;
        float* rec = p.stat_rec ? p.stat_rec + ((int64_t)(blockIdx.x * 4 + q) * p.rec_per_warp) * kRingRecFloats : nullptr;
        auto flush = [&]() {
            float wv[16];
#pragma unroll
            for (int g8 = 0; g8 < 8; ++g8) { wv[2 * g8] = s1[g8]; wv[2 * g8 + 1] = s2[g8]; s1[g8] = 0.f; s2[g8] = 0.f; }
            warp_reduce16(wv, lane);
            float* r = rec + (int64_t)nrec * kRingRecFloats;
            if ((lane & 1) == 0) r[lane >> 1] = wv[0];
            if (lane == 1) r[16] = __int_as_float(cur_n);
            ++nrec;
        };
        for (int it = 0; it < n_steps; ++it) {
            const int buf = it & (kRingTmemBufs - 1);
            mbar_wait(&tfull_bar[buf], (it / kRingTmemBufs) & 1);
            tc_fence_after_sync();
#pragma unroll
            for (int g = 0; g < G; ++g) {
                const bool valid = row < p.total_rows && y >= 1 && y <= p.H && x >= 1 && x <= p.W;
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((buf * G + g) * BN);
                uint32_t raw[BN];
#pragma unroll
                for (int c0 = 0; c0 < BN; c0 += 16) tmem_ld16_nowait(taddr + c0, *reinterpret_cast<uint32_t(*)[16]>(&raw[c0]));
                tmem_ld_wait();
                if (g == G - 1) {                                   // accumulators are in registers: hand the buffer back
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty_bar[buf]);
                }
                if (rec) {                                          // valid rows of one warp always belong to ONE sample
                    const unsigned m = __ballot_sync(0xffffffffu, valid);
                    if (m) {
                        const int nv = __shfl_sync(0xffffffffu, n, __ffs(m) - 1);
                        if (nv != cur_n) {
                            if (cur_n >= 0) flush();
                            cur_n = nv;
                        }
                    }
                }
                if (valid) {
                    __nv_bfloat16* out = p.dst + (int64_t)row * 8;
#pragma unroll
                    for (int c0 = 0; c0 < BN; c0 += 8) {
                        float v[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            v[i] = __uint_as_float(raw[c0 + i]);
                            if (p.bias) v[i] += __ldg(p.bias + c0 + i);
                        }
                        if (rec) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const int g8 = (c0 + i) / GS;
                                s1[g8] += v[i];
                                s2[g8] = fmaf(v[i], v[i], s2[g8]);
                            }
                        }
                        uint4 o;
                        o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
                        o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
                        *reinterpret_cast<uint4*>(out + (int64_t)(c0 / 8) * p.csd) = o;
                    }
                }
                row += 128; x += adv_x; y += adv_y;
                if (x >= p.pw) { x -= p.pw; ++y; }
                if (y >= HP) { y -= HP; ++n; }
            }
        }
        if (rec) {
            if (cur_n >= 0) flush();
            for (int j = nrec; j < p.rec_per_warp; ++j)
                if (lane == 1) rec[(int64_t)j * kRingRecFloats + 16] = __int_as_float(-1);
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// stats[n][g] = (mean, rstd) from the per-warp records of conv_ring_kernel, summed in a fixed order (deterministic)
__global__ void __launch_bounds__(256) gn_finalize_ring_kernel(const float* __restrict__ rec, float* __restrict__ stats,
                                                               int rec_per_cta, int num_ctas, int64_t px_per_cta, int64_t S, double count) {
    __shared__ double s_acc[16][16];
    const int n = blockIdx.x;
    const int idx = threadIdx.x & 15, tl = threadIdx.x >> 4;
    // only the CTAs whose pixel range [c * px_per_cta, (c + 1) * px_per_cta) meets sample n can hold records of it (each record
    // still carries its sample index; a CTA's range usually straddles two samples)
    const int c_lo = (int)(((int64_t)n * S) / px_per_cta);
    const int c_hi = min(num_ctas - 1, (int)((((int64_t)n + 1) * S - 1) / px_per_cta));
    double acc = 0.0;
    for (int r = c_lo * rec_per_cta + tl; r < (c_hi + 1) * rec_per_cta; r += 16) {
        const float* e = rec + (int64_t)r * kRingRecFloats;
        if (__float_as_int(e[16]) == n) acc += (double)e[idx];
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

// fp32 NCHW [Cout][Cin][ks][ks] -> wf[tap][Cin/8][Cout][8] and wd[taps-1-tap][Cout/8][Cin][8]: the no-swizzle K-major
// core-matrix order of the B operand (8 output rows x 8 K elements = 128 contiguous bytes)
__global__ void pack_weights_ring_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wf,
                                         __nv_bfloat16* __restrict__ wd, int Cout, int Cin, int taps) {
    const int64_t total = (int64_t)Cout * Cin * taps;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int tap = (int)(i % taps);
        const int ci = (int)((i / taps) % Cin);
        const int co = (int)(i / ((int64_t)taps * Cin));
        const __nv_bfloat16 v = __float2bfloat16_rn(w[i]);
        wf[(((int64_t)tap * (Cin / 8) + ci / 8) * Cout + co) * 8 + (ci & 7)] = v;
        if (wd) wd[(((int64_t)(taps - 1 - tap) * (Cout / 8) + co / 8) * Cin + ci) * 8 + (co & 7)] = v;
    }
}

struct RingPlan { int G, R, halo_lo, halo_hi, per_sm; uint32_t w_bytes, smem_bytes; bool ok; };

// largest block (G = 2, then 1) and deepest look-ahead that fit shared memory
// MRISR_RING_CTAS=1 keeps one CTA per SM (A/B measurements); default 2 where shared memory and TMEM allow
static int ring_ctas_wanted() {
    static const int want = [] { const char* e = getenv("MRISR_RING_CTAS"); return e ? atoi(e) : 2; }();
    return want;
}
constexpr uint32_t kRingSmemTwoCtas = 113u * 1024u;      // (228 KB - 2 x 1 KB reserved) / 2

static RingPlan ring_plan(int W, int Cin, int Cout) {
    RingPlan pl;
    const int pw = W + 2;
    pl.w_bytes = 9u * Cin * Cout * 2u;
    const uint32_t w_pad = (pl.w_bytes + 1023u) & ~1023u;
    pl.ok = false; pl.R = 0; pl.G = 0; pl.smem_bytes = 0; pl.halo_lo = pl.halo_hi = 0; pl.per_sm = 1;
    // pass 0: a ring small enough for two CTAs per SM (measured 1.6x on the 16-channel 512^2 layers: one CTA's epilogue and
    // producer stalls hide behind the other's MMAs); pass 1: the deepest ring that fits one CTA
    for (int pass = ring_ctas_wanted() >= 2 ? 0 : 1; pass < 2 && !pl.ok; ++pass) {
        const uint32_t limit = pass == 0 ? kRingSmemTwoCtas : 225u * 1024u;
        for (int G = 2; G >= 1 && !pl.ok; --G) {
            if (pass == 0 && 4 * G * Cout > 256) continue;              // TMEM columns of two CTAs
            const int BLK = 128 * G;
            const int lo = (pw + 1 + BLK - 1) / BLK, hi = (BLK + pw) / BLK;
            if ((int64_t)lo * BLK > planar_guard(W)) continue;
            const int live = lo + hi + 1;
            for (int ahead = 4; ahead >= (pass == 0 ? 2 : 1); --ahead) {      // two CTAs with look-ahead 1 measured slower than one
                const int R = live + ahead;
                const uint32_t need = w_pad + (uint32_t)(Cin / 8) * ((uint32_t)R * 2048u * G + 2048u) + 1024u;
                if (R <= kRingMaxSlots && need <= limit) {
                    pl.G = G; pl.R = R; pl.halo_lo = lo; pl.halo_hi = hi; pl.smem_bytes = need; pl.ok = true;
                    pl.per_sm = pass == 0 ? 2 : 1;
                    break;
                }
            }
        }
    }
    return pl;
}

template <int BN, int KP, int G>
static int launch_ring_cfg(const ConvRingParams& p, int grid, uint32_t smem, cudaStream_t s) {
    static PerDeviceFlag attr_done_pd; bool& attr_done = attr_done_pd.cur();
    if (!attr_done) {
        MRISR_CUDA_CHECK(cudaFuncSetAttribute(conv_ring_kernel<BN, KP, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        attr_done = true;
    }
    conv_ring_kernel<BN, KP, G><<<grid, 192, smem, s>>>(p);
    MRISR_LAUNCH_CHECK();
    return 0;
}
template <int BN, int KP>
static int launch_ring_g(int G, const ConvRingParams& p, int grid, uint32_t smem, cudaStream_t s) {
    return G == 2 ? launch_ring_cfg<BN, KP, 2>(p, grid, smem, s) : launch_ring_cfg<BN, KP, 1>(p, grid, smem, s);
}
template <int BN>
static int launch_ring_kp(int KP, int G, const ConvRingParams& p, int grid, uint32_t smem, cudaStream_t s) {
    switch (KP) {
        case 2: return launch_ring_g<BN, 2>(G, p, grid, smem, s);
        case 4: return launch_ring_g<BN, 4>(G, p, grid, smem, s);
        default: return launch_ring_g<BN, 8>(G, p, grid, smem, s);
    }
}

}  // namespace tc

using bf16 = __nv_bfloat16;

bool conv_ring_supported(int H, int W, int C1, int C2, int Cout, int ks) {
    const int Cin = C1 + C2;
    if (!(Cout == 16 || Cout == 32 || Cout == 64)) return false;
    if (C1 < 8 || C1 % 8 || C2 % 8 || !(Cin == 16 || Cin == 32 || Cin == 64) || ks != 3) return false;
    if (W < 16 || H < 16) return false;      // the epilogue relies on >= 32 dead pixels between the valid pixels of two samples
    return tc::ring_plan(W, Cin, Cout).ok;
}

// geometry shared by launch_conv_ring and the statistics finalize
struct RingGeom { int G, num_steps, steps_per_cta, grid, rec_per_warp; };
static RingGeom ring_geom(int N, int H, int W, int Cin, int Cout) {
    const tc::RingPlan pl = tc::ring_plan(W, Cin, Cout);
    RingGeom g;
    g.G = pl.G;
    const int64_t rows = (int64_t)N * (H + 2) * (W + 2);
    const int BLK = 128 * pl.G;
    g.num_steps = (int)((rows + BLK - 1) / BLK);
    // two CTAs per SM when shared memory and TMEM (4 accumulator buffers x G x Cout columns each) allow: one CTA's epilogue and
    // producer stalls overlap the other's MMAs (MRISR_RING_CTAS=1 disables)
    g.steps_per_cta = ceil_div(g.num_steps, tc::num_sms() * pl.per_sm);
    g.grid = ceil_div(g.num_steps, g.steps_per_cta);
    g.rec_per_warp = (int)(((int64_t)g.steps_per_cta * BLK) / ((int64_t)(H + 2) * (W + 2))) + 2;
    return g;
}

int conv_ring_stat_elems(int N, int H, int W, int Cin, int Cout) {
    if (Cin % 16 || Cout % 16 || !tc::ring_plan(W, Cin, Cout).ok) return 0;
    const RingGeom g = ring_geom(N, H, W, Cin, Cout);
    return g.grid * 4 * g.rec_per_warp * tc::kRingRecFloats;
}

int launch_gn_finalize_ring(const float* rec, float* stats, int N, int H, int W, int Cin, int Cout, double count, cudaStream_t s) {
    ProfScope prof_("gn_finalize", s);
    const RingGeom g = ring_geom(N, H, W, Cin, Cout);
    tc::gn_finalize_ring_kernel<<<N, 256, 0, s>>>(rec, stats, 4 * g.rec_per_warp, g.grid, (int64_t)g.steps_per_cta * 128 * g.G,
                                                  (int64_t)(H + 2) * (W + 2), count);
    MRISR_LAUNCH_CHECK();
    return 0;
}

int launch_conv_ring(const ConvArgs<bf16>& a, cudaStream_t s) {
    const int C1 = a.src1.C, C2 = a.src2.p ? a.src2.C : 0, Cin = C1 + C2, Cout = a.dst.C;
    const int H = a.dst.H, W = a.dst.W, N = a.dst.N;
    if (!conv_ring_supported(H, W, C1, C2, Cout, a.ks) || !a.src1.planar() || (C2 && !a.src2.planar()) || !a.dst.planar()) {
        set_error("conv_ring: unsupported shape/layout C1=%d C2=%d Cout=%d ks=%d %dx%d", C1, C2, Cout, a.ks, H, W);
        return -2;
    }
    const char* tag = "conv_ring";
    char tbuf[96];
    if (prof_on()) {
        snprintf(tbuf, sizeof(tbuf), "conv_ring:k%d:%d>%d@%dx%d", a.ks, Cin, Cout, H, W);
        tag = prof_intern(tbuf);
    }
    ProfScope prof_(tag, s);
    const tc::RingPlan pl = tc::ring_plan(W, Cin, Cout);
    const RingGeom g = ring_geom(N, H, W, Cin, Cout);
    tc::ConvRingParams p;
    p.src1 = a.src1.p; p.cs1 = a.src1.cs; p.planes1 = C1 / 8;
    p.src2 = C2 ? a.src2.p : a.src1.p; p.cs2 = C2 ? a.src2.cs : 0;
    p.w = a.w; p.dst = a.dst.p; p.csd = a.dst.cs; p.bias = a.bias;
    p.stat_rec = a.stat_part; p.rec_per_warp = g.rec_per_warp;
    p.pw = W + 2; p.S = (H + 2) * (W + 2); p.H = H; p.W = W;
    p.total_rows = (int)((int64_t)N * p.S);
    p.num_steps = g.num_steps; p.steps_per_cta = g.steps_per_cta;
    p.R = pl.R; p.halo_lo = pl.halo_lo; p.halo_hi = pl.halo_hi; p.w_bytes = pl.w_bytes;
    for (int t = 0; t < 9; ++t) p.sh[t] = pl.halo_lo * 128 * pl.G + (t / 3 - 1) * p.pw + (t % 3 - 1);
    switch (Cout) {
        case 16: return tc::launch_ring_kp<16>(Cin / 8, pl.G, p, g.grid, pl.smem_bytes, s);
        case 32: return tc::launch_ring_kp<32>(Cin / 8, pl.G, p, g.grid, pl.smem_bytes, s);
        default: return tc::launch_ring_kp<64>(Cin / 8, pl.G, p, g.grid, pl.smem_bytes, s);
    }
}

int launch_pack_weights_ring(const float* w, bf16* wf, bf16* wd, int Cout, int Cin, int ks, cudaStream_t s) {
    ProfScope prof_("pack_weights", s);
    const int64_t total = (int64_t)Cout * Cin * ks * ks;
    int blocks = (int)ceil_div64(total, 256);
    if (blocks > 1184) blocks = 1184;
    tc::pack_weights_ring_kernel<<<blocks, 256, 0, s>>>(w, wf, wd, Cout, Cin, ks * ks);
    MRISR_LAUNCH_CHECK();
    return 0;
}

}  // namespace mrisr
