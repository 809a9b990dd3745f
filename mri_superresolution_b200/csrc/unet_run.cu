// Forward and backward schedules of UNetSuperRes over the workspace (see unet_plan.cuh).
#include <stdlib.h>
#include "../../include/mrisr.h"
#include "unet_plan.cuh"

namespace mrisr {

#define TRY(expr) do { int _rc = (expr); if (_rc) return _rc; } while (0)

static inline bool is_up1x1(int l) { return l == 8 || l == 11 || l == 14; }

// Arena offset from which every gradient is final once the backward schedule has finished conv block `l`
// (schedule order 19, 17, 18, 16, 15, ..., 0; alpha, offset 0, is produced with block 19).
static int64_t grads_final_from(const ParamLayout& pl, const ConvDesc (&conv)[kNumConv], int l) {
    if (l >= 19 || l == 17) return pl.off[conv[19].w];
    if (l == 18) return pl.off[conv[17].w];
    return l <= 0 ? 0 : pl.off[conv[l].w];
}

// Side stream (default on, MRISR_BWD_SIDE=0 disables; 0.2 ms of the step on B200): weight-gradient kernels run on a second stream,
// forked from the main stream once their layer's dR is ready and joined before a gradient-readiness mark is recorded / at the
// end of backward.  They only read activations and dR (never rewritten during backward: every tensor of the workspace has its
// own buffer) and write the gradient arena and their own partial-sum scratch, so they are independent of the data-gradient
// chain; the chain is the critical path and its small deep-level kernels leave SMs idle that the weight gradients can use.
// All 20 convolutions' operand copies in ONE launch (bf16 mode): fp32 [Cout][Cin][ks][ks] -> the forward and data-gradient
// layouts of the layer's engine (conv_tc.cu / conv_ring.cu / conv_generic.cu state them).  20 separate launches of 5 - 6 us
// each were 0.16 ms of the step for 7 MB of traffic.
struct PackEntry { const float* w; __nv_bfloat16* wf; __nv_bfloat16* wd; int Cout, Cin, taps, kind; long long first; };
struct PackTable { PackEntry e[kNumConv]; int n; long long total; };
__global__ void __launch_bounds__(256) pack_weights_all_kernel(const __grid_constant__ PackTable tb) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < tb.total; i += (long long)gridDim.x * blockDim.x) {
        int l = 0;
        while (l + 1 < tb.n && i >= tb.e[l + 1].first) ++l;
        const PackEntry& p = tb.e[l];
        const long long j = i - p.first;
        const int taps = p.taps, Cin = p.Cin, Cout = p.Cout;
        const int tap = (int)(j % taps);
        const int ci = (int)((j / taps) % Cin);
        const int co = (int)(j / ((long long)taps * Cin));
        const __nv_bfloat16 v = __float2bfloat16_rn(p.w[j]);
        if (p.kind == ENG_TC) {
            p.wf[((long long)tap * Cout + co) * Cin + ci] = v;
            if (p.wd) p.wd[((long long)(taps - 1 - tap) * Cin + ci) * Cout + co] = v;
        } else if (p.kind == ENG_RING) {
            p.wf[(((long long)tap * (Cin / 8) + ci / 8) * Cout + co) * 8 + (ci & 7)] = v;
            if (p.wd) p.wd[(((long long)(taps - 1 - tap) * (Cout / 8) + co / 8) * Cin + ci) * 8 + (co & 7)] = v;
        } else {
            p.wf[((long long)tap * Cin + ci) * Cout + co] = v;
            if (p.wd) p.wd[((long long)(taps - 1 - tap) * Cout + co) * Cin + ci] = v;
        }
    }
}

// Created on first use outside stream capture (the graph helpers warm up eagerly first); fork / join by events, which a
// capturing stream records as graph dependencies.
struct SideStream {
    cudaStream_t stream = nullptr;
    cudaEvent_t fork_ev[kNumConv] = {};
    cudaEvent_t join_ev = nullptr;
    bool ok = false;
};
static SideStream* side_stream() {
    // on by default: measured on B200 in round 2 under graph replay, 17.60 -> 17.41 ms per step (the small deep-level kernels of
    // the data-gradient chain leave SMs idle that the weight-gradient kernels fill); MRISR_BWD_SIDE=0 keeps one stream
    static const bool enabled = [] { const char* e = getenv("MRISR_BWD_SIDE"); return !(e && e[0] == '0'); }();
    if (!enabled) return nullptr;
    if (prof_on()) return nullptr;        // per-kernel attribution (mrisr_prof_*): one stream, so a kernel's events time it alone
    static SideStream ss;
    static bool tried = false;
    if (!tried) {
        tried = true;
        int lo = 0, hi = 0;
        bool good = cudaDeviceGetStreamPriorityRange(&lo, &hi) == cudaSuccess &&
                    cudaStreamCreateWithPriority(&ss.stream, cudaStreamNonBlocking, lo) == cudaSuccess &&
                    cudaEventCreateWithFlags(&ss.join_ev, cudaEventDisableTiming) == cudaSuccess;
        for (int l = 0; good && l < kNumConv; ++l) good = cudaEventCreateWithFlags(&ss.fork_ev[l], cudaEventDisableTiming) == cudaSuccess;
        ss.ok = good;
        if (!good) cudaGetLastError();
    }
    return ss.ok ? &ss : nullptr;
}

template <typename T>
struct Runner {
    Workspace<T>& ws;
    const ParamLayout& pl;
    const float* params;
    cudaStream_t s;
    const float* P(int idx) const { return params + pl.off[idx]; }
    int eng(int l) const { return sizeof(T) == 2 ? layer_engine(ws.conv[l], l, ws.H, ws.W, ws.f) : (int)ENG_GENERIC; }

    // conv l: (src1 [, src2]) -> dst, GroupNorm partial statistics when `stats`
    int conv(int l, Act<T> src1, Act<T> src2, Act<T> dst, bool want_stats) {
        const ConvDesc& c = ws.conv[l];
        ConvArgs<T> a;
        a.src1 = src1; a.src2 = src2; a.w = ws.wf[l];
        a.bias = c.bias >= 0 ? P(c.bias) : nullptr;
        a.dst = dst; a.stat_part = want_stats ? ws.stat_part : nullptr; a.ks = c.ks;
        if constexpr (sizeof(T) == 2) {
            if (l == 0 && conv_first_supported(src1, dst, c.ks)) {
                TRY(launch_conv_first(a, s));
                if (want_stats) TRY(launch_gn_finalize(ws.stat_part, ws.stats[l], ws.N, conv_first_stat_slots(dst.H, dst.W, dst.C),
                                                       (double)(dst.C / kGroups) * dst.H * dst.W, s));
                return 0;
            }
            if (eng(l) != ENG_GENERIC) {
                const double count = (double)(dst.C / kGroups) * dst.H * dst.W;
                if (eng(l) == ENG_RING) {
                    TRY(launch_conv_ring(a, s));
                    if (want_stats) TRY(launch_gn_finalize_ring(ws.stat_part, ws.stats[l], ws.N, dst.H, dst.W, c.cin, c.cout, count, s));
                } else if (dst.C > 256) {
                    // two column blocks per tile: the statistics epilogue covers one block's groups only -> separate pass
                    // (512 channels only occur at the 32 x 32 bottleneck of base_filters = 64: 1 MB per sample)
                    a.stat_part = nullptr;
                    TRY(launch_conv_tc(a, s));
                    if (want_stats) {
                        TRY(launch_gn_stats<T>(dst, ws.stat_part, s));
                        TRY(launch_gn_finalize(ws.stat_part, ws.stats[l], ws.N, stat_slots_generic(dst.H, dst.W, dst.C), count, s));
                    }
                } else {
                    TRY(launch_conv_tc(a, s));
                    if (want_stats) TRY(launch_gn_finalize_tc(ws.stat_part, ws.stats[l], ws.N, dst.H, dst.W, count, s));
                }
                return 0;
            }
        }
        TRY(launch_conv_generic<T>(a, s));
        if (want_stats) {
            const int slots = conv_generic_stat_slots(dst.H, dst.W, dst.C);
            TRY(launch_gn_finalize(ws.stat_part, ws.stats[l], ws.N, slots,
                                   (double)(dst.C / kGroups) * dst.H * dst.W, s));
        }
        return 0;
    }
    int gn_act(int l, Act<T> pooled) {
        const ConvDesc& c = ws.conv[l];
        return launch_gn_apply<T>(ws.R[l], ws.stats[l], P(c.gamma), P(c.beta), ws.A[l], pooled, s);
    }
    // conv -> GN -> LReLU (-> pool)
    int block(int l, Act<T> src1, Act<T> src2, Act<T> pooled) {
        TRY(conv(l, src1, src2, ws.R[l], true));
        return gn_act(l, pooled);
    }
    // Up.up: bilinear x2 -> conv1x1 -> GN -> LReLU, evaluated as conv1x1 at low resolution followed by the
    // interpolation (both are linear and the conv has no bias, models/unet_model.py:70-72)
    int up_branch(int l, Act<T> src) {
        const Act<T> none{nullptr, 0, 0, 0, 0};
        TRY(conv(l, src, none, ws.Q[l], false));
        // the interpolation kernel sums what it stores: no separate statistics pass over the up-sampled tensor
        int slots = 0;
        const int rc = launch_upsample2x_stats<T>(ws.Q[l], ws.R[l], ws.stat_part, ws.stat_part_floats, &slots, s);
        if (rc != 0 && rc != MRISR_NOT_APPLICABLE) return rc;
        if (rc == MRISR_NOT_APPLICABLE) {
            TRY(launch_upsample2x<T>(ws.Q[l], ws.R[l], s));
            TRY(launch_gn_stats<T>(ws.R[l], ws.stat_part, s));
            slots = stat_slots_generic(ws.R[l].H, ws.R[l].W, ws.R[l].C);
        }
        TRY(launch_gn_finalize(ws.stat_part, ws.stats[l], ws.N, slots, (double)(ws.R[l].C / kGroups) * ws.R[l].H * ws.R[l].W, s));
        return gn_act(l, none);
    }
    BlendArgs<T> blend_args() {
        BlendArgs<T> b;
        b.r_bil = ws.R[17]; b.st_bil = ws.stats[17]; b.g_bil = P(ws.conv[17].gamma); b.b_bil = P(ws.conv[17].beta);
        b.r_ps = ws.R[18];  b.st_ps = ws.stats[18];  b.g_ps = P(ws.conv[18].gamma);  b.b_ps = P(ws.conv[18].beta);
        b.alpha = P(0); b.xm = ws.XM;
        return b;
    }

    int forward(const float* x, float* y, bool train) {
        const Act<T> none{nullptr, 0, 0, 0, 0};
        if constexpr (sizeof(T) == 2) {
            PackTable tb;
            tb.n = kNumConv;
            long long first = 0;
            for (int l = 0; l < kNumConv; ++l) {
                const ConvDesc& c = ws.conv[l];
                tb.e[l] = PackEntry{P(c.w), ws.wf[l], train && l > 0 ? ws.wd[l] : nullptr, c.cout, c.cin, c.ks * c.ks, eng(l), first};
                first += (long long)c.cout * c.cin * c.ks * c.ks;
            }
            tb.total = first;
            ProfScope prof_("pack_weights", s);
            pack_weights_all_kernel<<<(int)((first + 255) / 256 > 1184 ? 1184 : (first + 255) / 256), 256, 0, s>>>(tb);
            MRISR_LAUNCH_CHECK();
        }
        for (int l = 0; l < kNumConv && sizeof(T) != 2; ++l) {
            const ConvDesc& c = ws.conv[l];
            if constexpr (sizeof(T) == 2) {
                if (eng(l) != ENG_GENERIC) {
                    TRY((eng(l) == ENG_RING ? launch_pack_weights_ring : launch_pack_weights_tc)(
                        P(c.w), ws.wf[l], train && l > 0 ? ws.wd[l] : nullptr, c.cout, c.cin, c.ks, s));
                    continue;
                }
            }
            TRY(launch_pack_weights<T>(P(c.w), ws.wf[l], train && l > 0 ? ws.wd[l] : nullptr, c.cout, c.cin, c.ks, s));
        }
        TRY(launch_pack_input<T>(x, ws.X0, s));
        TRY(block(0, ws.X0, none, none));
        TRY(block(1, ws.A[0], none, ws.P[1]));
        TRY(block(2, ws.P[1], none, none));
        TRY(block(3, ws.A[2], none, ws.P[3]));
        TRY(block(4, ws.P[3], none, none));
        TRY(block(5, ws.A[4], none, ws.P[5]));
        TRY(block(6, ws.P[5], none, none));
        TRY(block(7, ws.A[6], none, none));
        TRY(up_branch(8, ws.A[7]));
        TRY(block(9, ws.A[5], ws.A[8], none));          // cat([x3, up]) models/unet_model.py:93
        TRY(block(10, ws.A[9], none, none));
        TRY(up_branch(11, ws.A[10]));
        TRY(block(12, ws.A[3], ws.A[11], none));
        TRY(block(13, ws.A[12], none, none));
        TRY(up_branch(14, ws.A[13]));
        TRY(block(15, ws.A[1], ws.A[14], none));
        TRY(block(16, ws.A[15], none, none));
        TRY(launch_upsample2x<T>(ws.A[16], ws.UB, s));  // final_up_bilinear.0
        TRY(conv(17, ws.UB, none, ws.R[17], true));
        TRY(conv(18, ws.A[16], none, ws.R[18], true));  // GN of the shuffled tensor == 8-group GN of the raw 2f channels
        TRY(launch_head_blend<T>(blend_args(), s));
        TRY(conv(19, ws.XM, none, ws.R[19], true));
        const ConvDesc& c19 = ws.conv[19];
        TRY(launch_final_head<T>(ws.R[19], ws.stats[19], P(c19.gamma), P(c19.beta), P(62), P(63), y, s));
        return 0;
    }

    // ---- backward helpers ----
    float* grads;
    // gradient-readiness marks (include/mrisr.h, mrisr_unet_backward_marked): event k is recorded on the stream as soon as
    // every arena element at offset >= mark_off[k] has its final value
    int n_marks = 0;
    const int64_t* mark_off = nullptr;
    void* const* mark_ev = nullptr;
    unsigned marks_done = 0;
    SideStream* side = nullptr;           // weight gradients on a second stream (MRISR_BWD_SIDE=1), else nullptr
    bool side_pending = false;
    int side_join() {
        if (side && side_pending) {
            MRISR_CUDA_CHECK(cudaEventRecord(side->join_ev, side->stream));
            MRISR_CUDA_CHECK(cudaStreamWaitEvent(s, side->join_ev, 0));
            side_pending = false;
        }
        return 0;
    }
    int block_done(int l) {
        const int64_t from = grads_final_from(pl, ws.conv, l);
        for (int k = 0; k < n_marks; ++k)
            if (!(marks_done >> k & 1u) && mark_off[k] >= from) {
                TRY(side_join());                               // the mark covers weight gradients still on the side stream
                MRISR_CUDA_CHECK(cudaEventRecord((cudaEvent_t)mark_ev[k], s));
                marks_done |= 1u << k;
            }
        return 0;
    }
    float* GP(int idx) const { return grads + pl.off[idx]; }

    GnBwdArgs<T> gn_args(int l, int kind, Act<T> g, int coff) {
        const ConvDesc& c = ws.conv[l];
        GnBwdArgs<T> a;
        memset(&a, 0, sizeof(a));
        a.kind = kind; a.r = ws.R[l]; a.stats = ws.stats[l]; a.gamma = P(c.gamma); a.beta = P(c.beta);
        a.g = g; a.coff = coff; a.part = ws.gn_part; a.mstat = ws.mstat;
        a.dgamma = GP(c.gamma); a.dbeta = GP(c.beta); a.dr = ws.dR[l]; a.alpha = P(0);
        return a;
    }
    // weight gradient + (optional) data gradient of conv l given dy
    int conv_bwd(int l, Act<T> src1, Act<T> src2, Act<T> dy, bool want_dgrad) {
        const ConvDesc& c = ws.conv[l];
        const Act<T> none{nullptr, 0, 0, 0, 0};
        WgradArgs<T> w;
        w.src1 = src1; w.src2 = src2; w.dy = dy; w.dw = GP(c.w); w.partial = ws.wg_part; w.ks = c.ks;
        const int e = eng(l);
        const bool use_tc = e != ENG_GENERIC;
        bool done = false;
        cudaStream_t sw = s;                                    // stream of the weight-gradient kernels
        if (side) {
            MRISR_CUDA_CHECK(cudaEventRecord(side->fork_ev[l], s));          // dy (and everything before it) is ready
            MRISR_CUDA_CHECK(cudaStreamWaitEvent(side->stream, side->fork_ev[l], 0));
            sw = side->stream;
            side_pending = true;
        }
        if constexpr (sizeof(T) == 2) {
            if (use_tc) {
                const int c2 = src2.p ? src2.C : 0;
                if (e == ENG_RING && wgrad_rows_supported(dy.N, dy.H, dy.W, src1.C, c2, dy.C, c.ks)) TRY(launch_wgrad_rows(w, sw));
                else if (e == ENG_RING) TRY(launch_wgrad_ring(w, sw));
                else if (wgrad_tc_supported(dy.H, dy.W, src1.C, c2, dy.C, c.ks)) TRY(launch_wgrad_tc(w, sw));
                else TRY(launch_wgrad_generic<T>(w, sw));
                done = true;
            }
            else if (l == 0 && conv_first_supported(src1, dy, c.ks)) { TRY(launch_wgrad_first(w, sw)); done = true; }
        }
        if (!done) TRY(launch_wgrad_generic<T>(w, sw));
        if (want_dgrad) {
            ConvArgs<T> a;
            a.src1 = dy; a.src2 = none; a.w = ws.wd[l]; a.bias = nullptr; a.dst = ws.G[l]; a.stat_part = nullptr; a.ks = c.ks;
            if constexpr (sizeof(T) == 2) {
                if (use_tc) TRY(e == ENG_RING ? launch_conv_ring(a, s) : launch_conv_tc(a, s));
            }
            if (!use_tc) TRY(launch_conv_generic<T>(a, s));
        }
        return 0;
    }
    // conv -> GN -> LReLU block: gradient source -> dR -> wgrad/dgrad
    int block_bwd(int l, int kind, Act<T> g, int coff, Act<T> gpool, Act<T> src1, Act<T> src2, bool want_dgrad = true,
                  float* third_row_sums = nullptr) {
        GnBwdArgs<T> a = gn_args(l, kind, g, coff);
        a.gpool = gpool;
        a.dwvec = third_row_sums;
        TRY(launch_gn_bwd<T>(a, s));
        return conv_bwd(l, src1, src2, ws.dR[l], want_dgrad);
    }
    int up_branch_bwd(int l, Act<T> g, int coff, Act<T> src) {
        const Act<T> none{nullptr, 0, 0, 0, 0};
        GnBwdArgs<T> a = gn_args(l, SRC_DIRECT, g, coff);
        TRY(launch_gn_bwd<T>(a, s));
        TRY(launch_upsample2x_bwd<T>(ws.dR[l], none, ws.dQ[l], s));
        return conv_bwd(l, src, none, ws.dQ[l], true);
    }

    int backward(const float* dy, const float* y, float* grads_) {
        grads = grads_;
        const Act<T> none{nullptr, 0, 0, 0, 0};
        const int f = ws.f;
        {   // final_conv.1..3 + sigmoid: rank-1 gradient source
            GnBwdArgs<T> a = gn_args(19, SRC_OUTER, none, 0);
            a.dy = dy; a.y = y; a.wvec = P(62); a.dwvec = GP(62); a.dbias = GP(63);
            TRY(launch_gn_bwd<T>(a, s));
            TRY(conv_bwd(19, ws.XM, none, ws.dR[19], true));          // G[19] = d(blend)
        }
        // the two blend-kind gn_bwd launches also leave  sum g * branch  per channel in ws.misc ([2][f/2]); d(alpha) is
        // formed from them before colsum reuses the scratch (no separate pass over both branches and the gradient)
        TRY(block_done(19));
        const int nch = ws.R[17].C;
        TRY(block_bwd(17, SRC_BLEND_BIL, ws.G[19], 0, none, ws.UB, none, true, ws.misc));
        TRY(block_bwd(18, SRC_BLEND_PS, ws.G[19], 0, none, ws.A[16], none, true, ws.misc + nch));
        TRY(launch_dalpha_from_channels(ws.misc, nch, P(0), GP(0), s));
        TRY(launch_colsum<T>(ws.dR[18], GP(ws.conv[18].bias), ws.misc, s));
        TRY(block_done(18));
        TRY(launch_upsample2x_bwd<T>(ws.G[17], ws.G[18], ws.DY3, s));
        TRY(block_bwd(16, SRC_DIRECT, ws.DY3, 0, none, ws.A[15], none));
        TRY(block_done(16));
        TRY(block_bwd(15, SRC_DIRECT, ws.G[16], 0, none, ws.A[1], ws.A[14]));     // G[15]: [skip x1 | up]
        TRY(block_done(15));
        TRY(up_branch_bwd(14, ws.G[15], f, ws.A[13]));
        TRY(block_done(14));
        TRY(block_bwd(13, SRC_DIRECT, ws.G[14], 0, none, ws.A[12], none));
        TRY(block_done(13));
        TRY(block_bwd(12, SRC_DIRECT, ws.G[13], 0, none, ws.A[3], ws.A[11]));
        TRY(block_done(12));
        TRY(up_branch_bwd(11, ws.G[12], 2 * f, ws.A[10]));
        TRY(block_done(11));
        TRY(block_bwd(10, SRC_DIRECT, ws.G[11], 0, none, ws.A[9], none));
        TRY(block_done(10));
        TRY(block_bwd(9, SRC_DIRECT, ws.G[10], 0, none, ws.A[5], ws.A[8]));
        TRY(block_done(9));
        TRY(up_branch_bwd(8, ws.G[9], 4 * f, ws.A[7]));
        TRY(block_done(8));
        TRY(block_bwd(7, SRC_DIRECT, ws.G[8], 0, none, ws.A[6], none));
        TRY(block_done(7));
        TRY(block_bwd(6, SRC_DIRECT, ws.G[7], 0, none, ws.P[5], none));           // G[6] = d(pool(x3))
        TRY(block_done(6));
        TRY(block_bwd(5, SRC_DIRECT_POOL, ws.G[9], 0, ws.G[6], ws.A[4], none));
        TRY(block_done(5));
        TRY(block_bwd(4, SRC_DIRECT, ws.G[5], 0, none, ws.P[3], none));
        TRY(block_done(4));
        TRY(block_bwd(3, SRC_DIRECT_POOL, ws.G[12], 0, ws.G[4], ws.A[2], none));
        TRY(block_done(3));
        TRY(block_bwd(2, SRC_DIRECT, ws.G[3], 0, none, ws.P[1], none));
        TRY(block_done(2));
        TRY(block_bwd(1, SRC_DIRECT_POOL, ws.G[15], 0, ws.G[2], ws.A[0], none));
        TRY(block_done(1));
        TRY(block_bwd(0, SRC_DIRECT, ws.G[1], 0, none, ws.X0, none, false));
        TRY(side_join());
        TRY(block_done(0));
        return 0;
    }
};

template <typename T>
static int forward_t(const float* params, const float* x, float* y, void* wsp, size_t wsb, int N, int H, int W, int f,
                     int train, cudaStream_t s) {
    Workspace<T> ws(wsp, N, H, W, f, train != 0);
    if (ws.bytes > wsb) { set_error("workspace too small: need %zu bytes, got %zu", ws.bytes, wsb); return MRISR_E_WORKSPACE; }
    ParamLayout pl(f);
    Runner<T> r{ws, pl, params, s};
    TRY(r.forward(x, y, train != 0));
    return 0;
}

template <typename T>
static int backward_t(const float* params, const float* dy, const float* y, float* grads, void* wsp, size_t wsb, int N,
                      int H, int W, int f, cudaStream_t s, int n_marks, const int64_t* mark_off, void* const* mark_ev) {
    Workspace<T> ws(wsp, N, H, W, f, true);
    if (ws.bytes > wsb) { set_error("workspace too small: need %zu bytes, got %zu", ws.bytes, wsb); return MRISR_E_WORKSPACE; }
    ParamLayout pl(f);
    Runner<T> r{ws, pl, params, s};
    r.n_marks = n_marks; r.mark_off = mark_off; r.mark_ev = mark_ev;
    r.side = side_stream();
    const int rc = r.backward(dy, y, grads);
    if (rc && r.side_pending) r.side_join();                    // never leave forked work unjoined (stream capture would fail to end)
    return rc;
}

template <typename T>
__global__ void unpack_act_kernel(Act<T> a, float* __restrict__ out) {
    const int64_t total = (int64_t)a.N * a.C * a.H * a.W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int x = (int)(i % a.W), y = (int)((i / a.W) % a.H);
        const int c = (int)((i / ((int64_t)a.W * a.H)) % a.C), n = (int)(i / ((int64_t)a.W * a.H * a.C));
        out[i] = to_f(*a.at(n, y, x, c));
    }
}

template <typename T>
static int debug_tensor_t(void* wsp, int N, int H, int W, int f, int kind, int l, float* out, int* dims, cudaStream_t s) {
    Workspace<T> ws(wsp, N, H, W, f, true);
    if (l < 0 || l >= kNumConv) { set_error("debug_tensor: bad layer %d", l); return MRISR_E_BADARG; }
    Act<T> a{nullptr, 0, 0, 0, 0};
    switch (kind) {
        case 0: a = ws.R[l]; break;   case 1: a = ws.A[l]; break;   case 2: a = ws.P[l]; break;
        case 3: a = ws.Q[l]; break;   case 4: a = ws.dR[l]; break;  case 5: a = ws.G[l]; break;
        case 6: a = ws.dQ[l]; break;  case 7: a = ws.UB; break;     case 8: a = ws.XM; break;
        case 9: a = ws.DY3; break;
        default: set_error("debug_tensor: bad kind %d", kind); return MRISR_E_BADARG;
    }
    if (!a.p) { set_error("debug_tensor: tensor (kind %d, layer %d) does not exist", kind, l); return MRISR_E_BADARG; }
    if (dims) { dims[0] = a.N; dims[1] = a.C; dims[2] = a.H; dims[3] = a.W; }
    if (out) {
        unpack_act_kernel<T><<<148 * 4, 256, 0, s>>>(a, out);
        MRISR_LAUNCH_CHECK();
    }
    return 0;
}

static int check_shape(int N, int H, int W, int f, int precision) {
    if (N <= 0 || H <= 0 || W <= 0) { set_error("N, H, W must be positive"); return MRISR_E_BADARG; }
    if (H % 8 || W % 8) { set_error("H and W must be multiples of 8 (three 2x2 poolings), got %dx%d", H, W); return MRISR_E_SHAPE; }
    if (f < 16 || f % 16) { set_error("base_filters must be a positive multiple of 16, got %d", f); return MRISR_E_SHAPE; }
    if (precision != MRISR_FP32 && precision != MRISR_BF16) { set_error("unknown precision %d", precision); return MRISR_E_DTYPE; }
    return 0;
}

}  // namespace mrisr

using namespace mrisr;

extern "C" {

int mrisr_version(void) { return MRISR_VERSION; }
const char* mrisr_last_error(void) { return g_err_ptr(); }

int mrisr_unet_param_tensors(int f) { return (int)ParamLayout(f).off.size(); }
int64_t mrisr_unet_param_elems(int f) { return ParamLayout(f).total; }
int64_t mrisr_unet_param_offset(int f, int i) { ParamLayout p(f); return i >= 0 && i < (int)p.off.size() ? p.off[i] : -1; }
int64_t mrisr_unet_param_size(int f, int i) { ParamLayout p(f); return i >= 0 && i < (int)p.size.size() ? p.size[i] : -1; }

size_t mrisr_unet_workspace_bytes(int N, int H, int W, int f, int precision, int train) {
    if (check_shape(N, H, W, f, precision)) return 0;
    if (precision == MRISR_FP32) return Workspace<float>(nullptr, N, H, W, f, train != 0).bytes;
    return Workspace<__nv_bfloat16>(nullptr, N, H, W, f, train != 0).bytes;
}

int mrisr_unet_forward(const float* params, const float* x, float* y, void* workspace, size_t workspace_bytes, int N,
                       int H, int W, int f, int precision, int train, void* stream) {
    if (!params || !x || !y || !workspace) { set_error("null pointer argument"); return MRISR_E_BADARG; }
    int rc = check_shape(N, H, W, f, precision);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (precision == MRISR_FP32) return forward_t<float>(params, x, y, workspace, workspace_bytes, N, H, W, f, train, s);
    return forward_t<__nv_bfloat16>(params, x, y, workspace, workspace_bytes, N, H, W, f, train, s);
}

int mrisr_unet_backward_marked(const float* params, const float* dy, const float* y, float* grads, float* dx,
                               void* workspace, size_t workspace_bytes, int N, int H, int W, int f, int precision,
                               void* stream, int n_marks, const int64_t* host_mark_offsets, void* const* host_mark_events) {
    if (!params || !dy || !y || !grads || !workspace) { set_error("null pointer argument"); return MRISR_E_BADARG; }
    if (dx) { set_error("gradient wrt the input image is not provided by this path (pass dx = NULL)"); return MRISR_E_BADARG; }
    if (n_marks < 0 || n_marks > 32 || (n_marks > 0 && (!host_mark_offsets || !host_mark_events))) {
        set_error("gradient marks: 0..32 marks with host arrays of offsets and events"); return MRISR_E_BADARG;
    }
    for (int k = 0; k < n_marks; ++k)
        if (host_mark_offsets[k] < 0 || !host_mark_events[k]) { set_error("gradient mark %d: negative offset or null event", k); return MRISR_E_BADARG; }
    int rc = check_shape(N, H, W, f, precision);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (precision == MRISR_FP32)
        return backward_t<float>(params, dy, y, grads, workspace, workspace_bytes, N, H, W, f, s, n_marks, host_mark_offsets, host_mark_events);
    return backward_t<__nv_bfloat16>(params, dy, y, grads, workspace, workspace_bytes, N, H, W, f, s, n_marks, host_mark_offsets, host_mark_events);
}

int mrisr_unet_backward(const float* params, const float* dy, const float* y, float* grads, float* dx, void* workspace,
                        size_t workspace_bytes, int N, int H, int W, int f, int precision, void* stream) {
    return mrisr_unet_backward_marked(params, dy, y, grads, dx, workspace, workspace_bytes, N, H, W, f, precision, stream,
                                      0, nullptr, nullptr);
}

int64_t mrisr_unet_grads_final_from(int f, int conv_index) {
    if (f < 16 || f % 16 || conv_index < 0 || conv_index >= kNumConv) return -1;
    ConvDesc conv[kNumConv];
    conv_table(f, conv);
    return grads_final_from(ParamLayout(f), conv, conv_index);
}

int mrisr_unet_layer_engine(int f, int conv_index, int H, int W) {
    if (f < 16 || f % 16 || conv_index < 0 || conv_index >= kNumConv || H <= 0 || W <= 0 || H % 8 || W % 8) return -1;
    ConvDesc conv[kNumConv];
    conv_table(f, conv);
    return layer_engine(conv[conv_index], conv_index, H, W, f);
}

void* mrisr_event_create(void) {
    cudaEvent_t e = nullptr;
    if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) { set_error("cudaEventCreate failed"); return nullptr; }
    return (void*)e;
}
int mrisr_event_destroy(void* event) { return event ? (int)cudaEventDestroy((cudaEvent_t)event) : 0; }
int mrisr_stream_wait_event(void* stream, void* event) {
    if (!event) { set_error("null event"); return MRISR_E_BADARG; }
    return (int)cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)event, 0);
}

int mrisr_unet_debug_tensor(void* workspace, int N, int H, int W, int f, int precision, int kind, int layer, float* out,
                            int* dims_host, void* stream) {
    if (!workspace) { set_error("null workspace"); return MRISR_E_BADARG; }
    int rc = check_shape(N, H, W, f, precision);
    if (rc) return rc;
    if (precision == MRISR_FP32) return debug_tensor_t<float>(workspace, N, H, W, f, kind, layer, out, dims_host, (cudaStream_t)stream);
    return debug_tensor_t<__nv_bfloat16>(workspace, N, H, W, f, kind, layer, out, dims_host, (cudaStream_t)stream);
}

size_t mrisr_l1_ssim_scratch_bytes(int N, int H, int W) { return l1_ssim_scratch_bytes(N, H, W); }

int mrisr_l1_ssim_fwd_bwd(const float* pred, const float* target, float* dpred, float* out5, void* scratch, int N,
                          int H, int W, float l1_weight, float ssim_weight, int window_size, float sigma,
                          float val_range, void* stream) {
    if (!pred || !target || !out5 || !scratch || N <= 0 || H <= 0 || W <= 0) { set_error("bad argument"); return MRISR_E_BADARG; }
    return launch_l1_ssim(pred, target, dpred, out5, scratch, N, H, W, l1_weight, ssim_weight, window_size, sigma,
                          val_range, (cudaStream_t)stream);
}

int mrisr_ssim_forward(const float* a, const float* b, float* per_sample, void* scratch, int N, int H, int W,
                       int window_size, float sigma, float val_range, void* stream) {
    if (!a || !b || !per_sample || !scratch || N <= 0 || H <= 0 || W <= 0) { set_error("bad argument"); return MRISR_E_BADARG; }
    return launch_ssim_forward(a, b, per_sample, scratch, N, H, W, window_size, sigma, val_range, (cudaStream_t)stream);
}

int mrisr_adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float b1, float b2, float eps,
                    float wd, int step, float grad_scale, void* stream) {
    if (!p || !g || !m || !v || n <= 0 || step < 1) { set_error("bad argument"); return MRISR_E_BADARG; }
    return launch_adam(p, g, m, v, n, lr, b1, b2, eps, wd, step, grad_scale, (cudaStream_t)stream);
}

int mrisr_adam_step_dev(float* p, const float* g, float* m, float* v, int64_t n, float b1, float b2, float eps, float wd,
                        void* state32, float grad_scale, void* stream) {
    if (!p || !g || !m || !v || !state32 || n <= 0) { set_error("bad argument"); return MRISR_E_BADARG; }
    return launch_adam_dev(p, g, m, v, n, b1, b2, eps, wd, state32, grad_scale, (cudaStream_t)stream);
}

}  // extern "C"
