// UNetSuperRes forward / backward schedule over the mrisr kernels (models/unet_model.py:189-211 and its
// autograd graph), plus the C ABI of include/mrisr.h.
#include <algorithm>
#include <stdarg.h>
#include <string.h>
#include "../../include/mrisr.h"
#include "unet_plan.cuh"

namespace mrisr {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

const char* g_err_ptr() { return g_err; }

ParamLayout::ParamLayout(int f) {
    auto add = [&](int64_t n) { off.push_back(total); size.push_back(n); total += n; };
    auto dconv = [&](int cin, int cout) {
        add((int64_t)cout * cin * 9); add(cout); add(cout);
        add((int64_t)cout * cout * 9); add(cout); add(cout);
    };
    add(1);                                   // alpha                      models/unet_model.py:163
    dconv(1, f); dconv(f, 2 * f); dconv(2 * f, 4 * f); dconv(4 * f, 8 * f);     // inc, down1..3   :137-140
    const int cup[3] = {8 * f, 4 * f, 2 * f}, cskip[3] = {4 * f, 2 * f, f}, co[3] = {4 * f, 2 * f, f};
    for (int i = 0; i < 3; ++i) {             // up1..3                     :144-146, :69-78
        add((int64_t)(cup[i] / 2) * cup[i]); add(cup[i] / 2); add(cup[i] / 2);
        dconv(cskip[i] + cup[i] / 2, co[i]);
    }
    add((int64_t)(f / 2) * f * 9); add(f / 2); add(f / 2);                      // final_up_bilinear     :150-155
    add((int64_t)(2 * f) * f * 9); add(2 * f); add(f / 2); add(f / 2);          // final_up_pixelshuffle :158,:101-103
    add((int64_t)(f / 2) * (f / 2) * 9); add(f / 2); add(f / 2); add(f / 2); add(1);   // final_conv     :166-173
}

void conv_table(int f, ConvDesc (&t)[kNumConv]) {
    const ConvDesc tab[kNumConv] = {
        {3, 1, f, 0, 1, -1, 2, 3},            {3, f, f, 0, 4, -1, 5, 6},
        {3, f, 2 * f, 1, 7, -1, 8, 9},        {3, 2 * f, 2 * f, 1, 10, -1, 11, 12},
        {3, 2 * f, 4 * f, 2, 13, -1, 14, 15}, {3, 4 * f, 4 * f, 2, 16, -1, 17, 18},
        {3, 4 * f, 8 * f, 3, 19, -1, 20, 21}, {3, 8 * f, 8 * f, 3, 22, -1, 23, 24},
        {1, 8 * f, 4 * f, 3, 25, -1, 26, 27}, {3, 8 * f, 4 * f, 2, 28, -1, 29, 30}, {3, 4 * f, 4 * f, 2, 31, -1, 32, 33},
        {1, 4 * f, 2 * f, 2, 34, -1, 35, 36}, {3, 4 * f, 2 * f, 1, 37, -1, 38, 39}, {3, 2 * f, 2 * f, 1, 40, -1, 41, 42},
        {1, 2 * f, f, 1, 43, -1, 44, 45},     {3, 2 * f, f, 0, 46, -1, 47, 48},     {3, f, f, 0, 49, -1, 50, 51},
        {3, f, f / 2, -1, 52, -1, 53, 54},    {3, f, 2 * f, 0, 55, 56, 57, 58},     {3, f / 2, f / 2, -1, 59, -1, 60, 61},
    };
    for (int i = 0; i < kNumConv; ++i) t[i] = tab[i];
}

template <typename T>
Act<T> Workspace<T>::make_act(Arena& ar, int lvl, int C) {
    Act<T> a;
    a.N = N; a.H = lvl_dim(H, lvl); a.W = lvl_dim(W, lvl); a.C = C;
    if (level_is_planar(f, lvl, H, W) && sizeof(T) == 2 && C >= 16 && C % 8 == 0) {
        // planar-8: C/8 planes of [guard | N*S pixels rounded to 128 | guard] x 8 channels.  The guards are only ever READ
        // (the ring kernels fetch whole 128-pixel blocks around a tile, conv_ring.cu); what they hold is irrelevant
        // because only border outputs depend on them and those are never stored.
        const int64_t guard = planar_guard(a.W);
        const int64_t plane = planar_plane(a.rows(), a.W);
        T* p = (T*)ar.take((size_t)plane * C * sizeof(T));
        a.ps = 8; a.cs = plane * 8;
        a.p = p ? p + guard * 8 : nullptr;
        return a;
    }
    const size_t guard = (((size_t)(a.W + 2) + 2 + 128) * C + 127) & ~(size_t)127;
    const size_t elems = (size_t)a.rows() * C + 2 * guard;
    T* p = (T*)ar.take(elems * sizeof(T));
    a.ps = C; a.cs = 8;
    a.p = p ? p + guard : nullptr;
    return a;
}

static inline bool is_up1x1(int l) { return l == 8 || l == 11 || l == 14; }

template <typename T>
Workspace<T>::Workspace(void* base, int N_, int H_, int W_, int f_, bool train) : N(N_), H(H_), W(W_), f(f_) {
    conv_table(f, conv);
    Arena ar(base);
    state = (int*)ar.take(256);
    X0 = make_act(ar, 0, 1);
    size_t part_max = 0, gnp_max = 0, wg_max = 0;
    for (int l = 0; l < kNumConv; ++l) {
        const ConvDesc& c = conv[l];
        const int rl = is_up1x1(l) ? c.lvl - 1 : c.lvl;            // level of the normalised tensor
        if (is_up1x1(l)) Q[l] = make_act(ar, c.lvl, c.cout); else Q[l] = Act<T>{nullptr, 0, 0, 0, 0};
        R[l] = make_act(ar, rl, c.cout);
        A[l] = (l == 17 || l == 18 || l == 19) ? Act<T>{nullptr, 0, 0, 0, 0} : make_act(ar, rl, c.cout);
        P[l] = (l == 1 || l == 3 || l == 5) ? make_act(ar, rl + 1, c.cout) : Act<T>{nullptr, 0, 0, 0, 0};
        stats[l] = (float*)ar.take((size_t)N * kGroups * 2 * sizeof(float));
        wf[l] = (T*)ar.take((size_t)c.cout * c.cin * c.ks * c.ks * sizeof(T));
        wd[l] = (T*)ar.take((size_t)c.cout * c.cin * c.ks * c.ks * sizeof(T));
        const int h = lvl_dim(H, rl), w = lvl_dim(W, rl);
        size_t slots = std::max(conv_generic_stat_slots(h, w, c.cout), stat_slots_generic(h, w, c.cout));
        part_max = std::max(part_max, (size_t)N * slots * kGroups * 2);
        part_max = std::max(part_max, (size_t)conv_tc_stat_elems(N, h, w));
        part_max = std::max(part_max, (size_t)conv_ring_stat_elems(N, h, w, c.cin, c.cout));
        if (l == 0) {
            part_max = std::max(part_max, (size_t)N * conv_first_stat_slots(h, w, c.cout) * kGroups * 2);
            wg_max = std::max(wg_max, wgrad_first_scratch_elems(N, h, w, c.cout));
        }
        gnp_max = std::max(gnp_max, (size_t)N * gn_bwd_slots(h, w, c.cout) * 3 * (c.cout + 8));
        wg_max = std::max(wg_max, wgrad_generic_scratch_elems(N, lvl_dim(H, c.lvl), lvl_dim(W, c.lvl), c.cin, c.cout, c.ks));
        if (sizeof(T) == 2 && layer_engine(c, l, H, W, f) == ENG_TC) {
            const int c1 = is_concat_conv(l) ? c.cin / 2 : c.cin;
            wg_max = std::max(wg_max, wgrad_tc_scratch_elems(N, lvl_dim(H, c.lvl), lvl_dim(W, c.lvl), c1, c.cin - c1, c.cout, c.ks));
        }
        if (sizeof(T) == 2 && layer_engine(c, l, H, W, f) == ENG_RING) {
            wg_max = std::max(wg_max, wgrad_ring_scratch_elems(N, lvl_dim(H, c.lvl), lvl_dim(W, c.lvl), c.cin, c.cout, c.ks));
            wg_max = std::max(wg_max, wgrad_rows_scratch_elems(N, lvl_dim(H, c.lvl), lvl_dim(W, c.lvl), c.cin, c.cout));
        }
    }
    UB = make_act(ar, -1, f);
    XM = make_act(ar, -1, f / 2);
    stat_part = (float*)ar.take(part_max * sizeof(float));
    stat_part_floats = part_max;
    if (train) {
        for (int l = 0; l < kNumConv; ++l) {
            const ConvDesc& c = conv[l];
            const int rl = is_up1x1(l) ? c.lvl - 1 : c.lvl;
            dR[l] = make_act(ar, rl, c.cout);
            dQ[l] = is_up1x1(l) ? make_act(ar, c.lvl, c.cout) : Act<T>{nullptr, 0, 0, 0, 0};
            G[l] = l == 0 ? Act<T>{nullptr, 0, 0, 0, 0} : make_act(ar, c.lvl, c.cin);
        }
        DY3 = make_act(ar, 0, f);
        gn_part = (float*)ar.take(gnp_max * sizeof(float));
        mstat = (float*)ar.take((size_t)N * kGroups * 2 * sizeof(float));
        wg_part = (float*)ar.take(wg_max * sizeof(float));
        const size_t misc_elems = std::max(colsum_scratch_elems(N, H, W, 2 * f), (size_t)4 * f);
        misc = (float*)ar.take(misc_elems * sizeof(float));
    } else {
        gn_part = mstat = wg_part = misc = nullptr;
    }
    bytes = ar.used;
}

template struct Workspace<float>;
template struct Workspace<__nv_bfloat16>;

}  // namespace mrisr
