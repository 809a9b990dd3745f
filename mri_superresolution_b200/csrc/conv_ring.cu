// tcgen05 implicit-GEMM 3x3 / 1x1 convolution for NARROW, HIGH-RESOLUTION layers (Cin <= 64 at 256^2 / 512^2), over the
// planar-8 bf16 layout (common.cuh: C/8 planes of [pixel][8 channels]).  Same math as conv_tc.cu (forward conv of
// models/unet_model.py:29,34,101,152,168 and, with tap-flipped transposed weights, its data gradient) but organised around
// what bounds these layers: with Cout <= 64 the tensor pipe is starved by operand traffic, and re-fetching the activation
// tile once per tap (9x from L2, one TMA request per 32..128-byte row) costs ~5x the HBM time of the layer.  Here
//   * a CTA walks a contiguous run of 128-pixel tiles and keeps a sliding window of the input in shared memory: a ring of
//     128-pixel blocks per 8-channel plane, each block ONE 2 KB cp.async.bulk copy, so every input byte is fetched once;
//   * the planar block IS the no-swizzle K-major core-matrix layout of tcgen05.mma (8 pixels x 16 B contiguous), so a tap
//     is just the same descriptor with its start address moved by shift * 16 B - no data movement per tap at all;
//     a mirror copy of ring slot 0 behind the last slot keeps every 128-pixel operand window contiguous across the wrap;
//   * the whole packed weight tensor (<= 72 KB) is resident in shared memory for the lifetime of the CTA.
// Warp roles as in conv_tc.cu: warp 0 producer (bulk copies), warp 1 TMEM allocator + MMA issuer, warps 2..5 epilogue.
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

constexpr int kRingMaxSlots = 16;

struct ConvRingParams {
    const __nv_bfloat16* src1; int64_t cs1; int planes1;    // planar-8 sources (pixel 0 of plane 0, plane stride in elements)
    const __nv_bfloat16* src2; int64_t cs2; int planes2;
    const __nv_bfloat16* w;         // packed [tap][Cin/8][Cout][8]
    __nv_bfloat16* dst; int64_t csd;                         // planar-8 output, Cout = BN
    const float* bias;
    float* stat_part;               // [num_tiles][2][8][2] or nullptr
    int taps, pw, S, H, W, total_rows, num_tiles, tiles_per_cta;
    int R;                          // ring slots (excluding the mirror slot)
    int halo_lo, halo_hi;           // blocks below / above the tile's own block that its window touches
    uint32_t w_bytes;
};

template <int BN>
__global__ void __launch_bounds__(192, 1) conv_ring_kernel(const ConvRingParams p) {
    constexpr uint32_t IDESC = make_idesc_bf16(128, BN, 0, 0);
    constexpr uint32_t TMEM_COLS = 2 * BN <= 32 ? 32 : 2 * BN <= 64 ? 64 : 128;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[kRingMaxSlots];
    __shared__ __align__(8) uint64_t empty_bar[kRingMaxSlots];
    __shared__ __align__(8) uint64_t tfull_bar[2];
    __shared__ __align__(8) uint64_t tempty_bar[2];
    __shared__ __align__(8) uint64_t w_bar;
    __shared__ uint32_t s_tmem;
    __shared__ float s_part[2][4][2][16];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KP = p.planes1 + p.planes2;                          // 8-channel planes of the input
    const uint32_t PLS = (uint32_t)(p.R + 1) * 2048u;              // plane stride inside the ring (bytes)
    uint8_t* w_smem = smem;
    uint8_t* ring = smem + ((p.w_bytes + 1023u) & ~1023u);
    const int t_begin = blockIdx.x * p.tiles_per_cta;
    const int t_end = min(p.num_tiles, t_begin + p.tiles_per_cta);
    const int b0 = t_begin - p.halo_lo;                            // first block this CTA ever loads (may be negative)
    const int num_blocks = (t_end - 1 + p.halo_hi) - b0 + 1;

    if (threadIdx.x == 0) {
        for (int i = 0; i < p.R; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 128); }
        mbar_init(&w_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(&s_tmem, TMEM_COLS);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = s_tmem;

    if (warp == 0) {
        // ===== producer: weights once, then one 2 KB bulk copy per (block, plane) =====
        if (lane == 0 && t_begin < t_end) {
            mbar_arrive_expect_tx(&w_bar, p.w_bytes);
            for (uint32_t off = 0; off < p.w_bytes; off += 16384u) {
                const uint32_t n = min(16384u, p.w_bytes - off);
                bulk_load(w_smem + off, reinterpret_cast<const uint8_t*>(p.w) + off, n, &w_bar);
            }
            for (int j = 0; j < num_blocks; ++j) {
                const int slot = j % p.R;
                mbar_wait(&empty_bar[slot], ((j / p.R) & 1) ^ 1);
                const bool mirror = slot == 0;
                mbar_arrive_expect_tx(&full_bar[slot], (uint32_t)KP * 2048u * (mirror ? 2u : 1u));
                const int64_t pix0 = (int64_t)(b0 + j) * 128;
                for (int k = 0; k < KP; ++k) {
                    const __nv_bfloat16* src = k < p.planes1 ? p.src1 + (int64_t)k * p.cs1 + pix0 * 8
                                                             : p.src2 + (int64_t)(k - p.planes1) * p.cs2 + pix0 * 8;
                    uint8_t* dst = ring + (uint32_t)k * PLS + (uint32_t)slot * 2048u;
                    bulk_load(dst, src, 2048u, &full_bar[slot]);
                    if (mirror) bulk_load(dst + (uint32_t)p.R * 2048u, src, 2048u, &full_bar[slot]);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0 && t_begin < t_end) {
            mbar_wait(&w_bar, 0);
            const uint32_t ring_addr = smem_u32(ring), w_addr = smem_u32(w_smem);
            const uint32_t w_lbo = (uint32_t)BN * 16u;             // next 8-channel K chunk of the weights
            const uint32_t w_tap = (uint32_t)KP * w_lbo;
            int ready = 0;                                         // blocks [0, ready) have landed
            int released = 0;                                      // blocks [0, released) handed back to the producer
            int it = 0;
            for (int tile = t_begin; tile < t_end; ++tile, ++it) {
                const int need = tile + p.halo_hi - b0;            // highest block this tile's window touches
                while (ready <= need) {
                    mbar_wait(&full_bar[ready % p.R], (ready / p.R) & 1);
                    ++ready;
                }
                const int buf = it & 1;
                mbar_wait(&tempty_bar[buf], ((it >> 1) & 1) ^ 1);
                tc_fence_after_sync();
                const uint32_t d = tmem_base + (uint32_t)(buf * BN);
                uint32_t acc = 0;
                for (int tap = 0; tap < p.taps; ++tap) {
                    const int shift = p.taps == 9 ? (tap / 3 - 1) * p.pw + (tap % 3 - 1) : 0;
                    const int P0 = tile * 128 + shift - b0 * 128;  // >= 0: window start relative to the first loaded pixel
                    const int jb = P0 >> 7, o = P0 & 127;
                    const uint32_t a0 = ring_addr + (uint32_t)(jb % p.R) * 2048u + (uint32_t)o * 16u;
                    const uint32_t w0 = w_addr + (uint32_t)tap * w_tap;
                    for (int kc = 0; kc < KP; kc += 2) {
                        const uint64_t adesc = make_smem_desc(a0 + (uint32_t)kc * PLS, PLS, 128u, 0u);
                        const uint64_t bdesc = make_smem_desc(w0 + (uint32_t)kc * w_lbo, w_lbo, 128u, 0u);
                        umma_bf16(d, adesc, bdesc, IDESC, acc);
                        acc = 1u;
                    }
                }
                umma_commit(&tfull_bar[buf]);
                // blocks below the next tile's window are dead once these MMAs have completed
                const int dead = (tile + 1 - p.halo_lo) - b0;
                for (; released < dead; ++released) umma_commit(&empty_bar[released % p.R]);
            }
        }
    } else {
        // ===== epilogue (TMEM lane quarter = warp % 4) =====
        const int q = warp & 3;
        constexpr int GS = BN / 8;
        int it = 0;
        for (int tile = t_begin; tile < t_end; ++tile, ++it) {
            const int buf = it & 1;
            mbar_wait(&tfull_bar[buf], (it >> 1) & 1);
            tc_fence_after_sync();
            const int row = tile * 128 + q * 32 + lane;
            const int n = row / p.S;
            const int rem = row - n * p.S;
            const int y = rem / p.pw, x = rem - y * p.pw;
            const bool valid = row < p.total_rows && y >= 1 && y <= p.H && x >= 1 && x <= p.W;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BN);
            __nv_bfloat16* out = p.dst + (int64_t)row * 8;
            float s1[8], s2[8];
#pragma unroll
            for (int g = 0; g < 8; ++g) s1[g] = s2[g] = 0.f;
#pragma unroll
            for (int c0 = 0; c0 < BN; c0 += 16) {
                float v[16];
                tmem_ld16(taddr + c0, v);
                if (p.bias) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] += __ldg(p.bias + c0 + i);
                }
                if (valid) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const int g = (c0 + i) / GS;
                        s1[g] += v[i];
                        s2[g] = fmaf(v[i], v[i], s2[g]);
                    }
                    uint4 o0, o1;
                    o0.x = pack_bf16x2(v[0], v[1]);   o0.y = pack_bf16x2(v[2], v[3]);
                    o0.z = pack_bf16x2(v[4], v[5]);   o0.w = pack_bf16x2(v[6], v[7]);
                    o1.x = pack_bf16x2(v[8], v[9]);   o1.y = pack_bf16x2(v[10], v[11]);
                    o1.z = pack_bf16x2(v[12], v[13]); o1.w = pack_bf16x2(v[14], v[15]);
                    *reinterpret_cast<uint4*>(out + (int64_t)(c0 / 8) * p.csd) = o0;
                    *reinterpret_cast<uint4*>(out + (int64_t)(c0 / 8 + 1) * p.csd) = o1;
                }
            }
            tc_fence_before_sync();
            mbar_arrive(&tempty_bar[buf]);
            if (p.stat_part) {
                const int n0 = (tile * 128) / p.S;
                const int slot = n - n0;
#pragma unroll
                for (int sl = 0; sl < 2; ++sl) {
                    const bool mine = valid && slot == sl;
                    if (__ballot_sync(0xffffffffu, mine) != 0u) {
                        float wv[16];
#pragma unroll
                        for (int g = 0; g < 8; ++g) { wv[2 * g] = mine ? s1[g] : 0.f; wv[2 * g + 1] = mine ? s2[g] : 0.f; }
                        warp_reduce16(wv, lane);
                        if ((lane & 1) == 0) s_part[it & 1][q][sl][lane >> 1] = wv[0];
                    } else if (lane < 16) {
                        s_part[it & 1][q][sl][lane] = 0.f;
                    }
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (q == 0) {
                    const int sl = lane >> 4, idx = lane & 15;
                    const float t = (s_part[it & 1][0][sl][idx] + s_part[it & 1][1][sl][idx]) +
                                    (s_part[it & 1][2][sl][idx] + s_part[it & 1][3][sl][idx]);
                    p.stat_part[((int64_t)tile * 2 + sl) * 16 + idx] = t;
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, TMEM_COLS);
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

struct RingPlan { int R, halo_lo, halo_hi; uint32_t w_bytes, smem_bytes; bool ok; };

static RingPlan ring_plan(int W, int Cin, int Cout, int ks) {
    RingPlan pl;
    const int pw = W + 2;
    pl.halo_lo = ks == 3 ? (pw + 1 + 127) / 128 : 0;
    pl.halo_hi = ks == 3 ? (pw + 128) / 128 : 0;
    const int live = pl.halo_lo + pl.halo_hi + 1;
    pl.w_bytes = (uint32_t)ks * ks * Cin * Cout * 2u;
    const uint32_t w_pad = (pl.w_bytes + 1023u) & ~1023u;
    pl.ok = false;
    pl.R = 0;
    pl.smem_bytes = 0;
    for (int ahead = 3; ahead >= 1; --ahead) {
        const int R = live + ahead;
        const uint32_t need = w_pad + (uint32_t)(Cin / 8) * (uint32_t)(R + 1) * 2048u + 1024u;
        if (R <= kRingMaxSlots && need <= 200u * 1024u) { pl.R = R; pl.smem_bytes = need; pl.ok = true; break; }
    }
    return pl;
}

}  // namespace tc

using bf16 = __nv_bfloat16;

bool conv_ring_supported(int H, int W, int C1, int C2, int Cout, int ks) {
    const int Cin = C1 + C2;
    if (!(Cout == 16 || Cout == 32 || Cout == 64)) return false;
    if (C1 < 8 || C1 % 8 || C2 % 8 || Cin % 16 || Cin > 64 || (ks != 1 && ks != 3)) return false;
    if ((int64_t)(H + 2) * (W + 2) < 128) return false;
    return tc::ring_plan(W, Cin, Cout, ks).ok;
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
    const tc::RingPlan pl = tc::ring_plan(W, Cin, Cout, a.ks);
    tc::ConvRingParams p;
    p.src1 = a.src1.p; p.cs1 = a.src1.cs; p.planes1 = C1 / 8;
    p.src2 = C2 ? a.src2.p : a.src1.p; p.cs2 = C2 ? a.src2.cs : 0; p.planes2 = C2 / 8;
    p.w = a.w; p.dst = a.dst.p; p.csd = a.dst.cs; p.bias = a.bias; p.stat_part = a.stat_part;
    p.taps = a.ks * a.ks; p.pw = W + 2; p.S = (H + 2) * (W + 2); p.H = H; p.W = W;
    const int64_t rows = (int64_t)N * p.S;
    p.total_rows = (int)rows; p.num_tiles = (int)((rows + 127) / 128);
    p.tiles_per_cta = ceil_div(p.num_tiles, tc::num_sms());
    p.R = pl.R; p.halo_lo = pl.halo_lo; p.halo_hi = pl.halo_hi; p.w_bytes = pl.w_bytes;
    const int grid = ceil_div(p.num_tiles, p.tiles_per_cta);
    static bool attr_done[3] = {false, false, false};
    auto launch = [&](auto kern, int idx) -> int {
        if (!attr_done[idx]) {
            MRISR_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 201 * 1024));
            attr_done[idx] = true;
        }
        kern<<<grid, 192, pl.smem_bytes, s>>>(p);
        MRISR_LAUNCH_CHECK();
        return 0;
    };
    switch (Cout) {
        case 16: return launch(tc::conv_ring_kernel<16>, 0);
        case 32: return launch(tc::conv_ring_kernel<32>, 1);
        default: return launch(tc::conv_ring_kernel<64>, 2);
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
