// Static description of UNetSuperRes (models/unet_model.py:129-175) and the workspace layout.
#pragma once
#include <stdlib.h>
#include <vector>
#include "kernels.cuh"

namespace mrisr {

constexpr int kNumConv = 20;     // GEMM-shaped convolutions (the final 1x1 -> 1 channel is fused into final_head)

// Parameter tensor indices in registration order (alpha first, then children in __init__ order).
struct ParamLayout {
    std::vector<int64_t> off, size;
    int64_t total = 0;
    explicit ParamLayout(int f);
};

// resolution level: 0 -> HxW, 1 -> H/2, 2 -> H/4, 3 -> H/8, -1 -> 2Hx2W
struct ConvDesc {
    int ks, cin, cout, lvl;
    int w, bias, gamma, beta;    // parameter indices (-1 if absent)
};
void conv_table(int f, ConvDesc (&t)[kNumConv]);

inline int lvl_dim(int d, int lvl) { return lvl < 0 ? 2 * d : d >> lvl; }
inline bool is_concat_conv(int l) { return l == 9 || l == 12 || l == 15; }   // DoubleConv fed by cat([skip, up])
// Resolution levels whose bf16 tensors use the planar-8 layout: full resolution and the 2x head, AS LONG AS the sliding-window
// kernels can hold three padded pixel rows of every 3x3 layer of that level in shared memory (conv_ring.cu: ring_plan).  At
// 512 x 512 input (BASELINE.json config 4) the 64-channel layer at 512^2 and the 32-channel layer at 1024^2 cannot; such a
// level falls back to NHWC + the TMA / tcgen05 kernels of conv_tc.cu as a whole (a tensor has ONE layout, and every layer of
// the level reads or writes it) instead of dropping single layers to the CUDA-core kernels (measured: 81 of 96 ms).
// MRISR_FORCE_NHWC=1 forces that fallback at every size (parity tests of the fallback at small sizes).
inline bool level_is_planar(int f, int lvl, int H, int W) {
    if (lvl > 0) return false;
    static const bool force_nhwc = getenv("MRISR_FORCE_NHWC") != nullptr;
    if (force_nhwc) return false;
    ConvDesc t[kNumConv];
    conv_table(f, t);
    const int h = lvl_dim(H, lvl), w = lvl_dim(W, lvl);
    for (int l = 0; l < kNumConv; ++l) {
        const ConvDesc& c = t[l];
        if (c.lvl != lvl || c.ks != 3 || c.cin < 16) continue;
        const int c1 = is_concat_conv(l) ? c.cin / 2 : c.cin, c2 = c.cin - c1;
        if (!(conv_ring_supported(h, w, c1, c2, c.cout, 3) && conv_ring_supported(h, w, c.cout, 0, c.cin, 3) &&
              wgrad_ring_supported(h, w, c1, c2, c.cout, 3)))
            return false;
    }
    return true;
}
// bf16 mode: does conv l (forward, data gradient and weight gradient) run on the tcgen05 kernels at input size HxW?
// ENG_TC: NHWC + TMA + swizzled tcgen05 (conv_tc.cu / wgrad_tc.cu), levels >= 1 (and wide levels <= 0, above).
// ENG_RING: planar-8 sliding-window tcgen05 (conv_ring.cu / wgrad_ring.cu), the narrow layers at levels <= 0.
enum Engine : int { ENG_GENERIC = 0, ENG_TC = 1, ENG_RING = 2 };
inline int layer_engine(const ConvDesc& c, int l, int H, int W, int f) {
    const int h = lvl_dim(H, c.lvl), w = lvl_dim(W, c.lvl);
    const int c1 = is_concat_conv(l) ? c.cin / 2 : c.cin, c2 = c.cin - c1;
    if (level_is_planar(f, c.lvl, H, W)) {
        const bool planar_ok = c1 >= 16 && (c2 == 0 || c2 >= 16) && c.cout >= 16;     // C = 8 tensors stay NHWC
        return planar_ok && conv_ring_supported(h, w, c1, c2, c.cout, c.ks) && conv_ring_supported(h, w, c.cout, 0, c.cin, c.ks) &&
                       wgrad_ring_supported(h, w, c1, c2, c.cout, c.ks) ? ENG_RING : ENG_GENERIC;
    }
    if ((int64_t)(h + 2) * (w + 2) < 128) return ENG_GENERIC;   // GroupNorm-statistics epilogue: <= 2 samples per tile
    if (!conv_tc_supported(h, w, c1, c2, c.cout, c.ks) || !conv_tc_supported(h, w, c.cout, 0, c.cin, c.ks)) return ENG_GENERIC;
    // 512-channel layers (base_filters = 64, the default of the reference's scripts/infer.py:473): conv_tc in two 256-column
    // blocks with the GroupNorm statistics from a separate pass (unet_run.cu), weight gradient over the two halves of dY
    return wgrad_tc_supported(h, w, c1, c2, c.cout, c.ks) ? ENG_TC : ENG_GENERIC;
}

// Bump allocator over the caller's workspace (base == nullptr -> size query only).
struct Arena {
    char* base;
    size_t used = 0;
    explicit Arena(void* b) : base((char*)b) {}
    void* take(size_t bytes) {
        void* p = base ? base + used : nullptr;
        used += (bytes + 255) & ~(size_t)255;
        return p;
    }
};

template <typename T>
struct Workspace {
    int N, H, W, f;
    ConvDesc conv[kNumConv];
    // forward tensors
    Act<T> X0, R[kNumConv], Q[kNumConv];          // Q: low-res 1x1 outputs of layers 8, 11, 14
    Act<T> A[kNumConv];                            // post GN+LReLU activation of conv l (l = 17, 18 unused)
    Act<T> P[kNumConv];                            // pooled activation (l = 1, 3, 5)
    Act<T> UB, XM;
    float* stats[kNumConv];                        // [N][8][2] mean, rstd
    float* stat_part;                              // shared partial-statistics scratch
    T* wf[kNumConv];                               // packed forward weights  [tap][Cin][Cout]
    T* wd[kNumConv];                               // packed dgrad weights    [tap'][Cout][Cin]
    // backward tensors (train only)
    Act<T> dR[kNumConv], dQ[kNumConv], G[kNumConv], DY3;
    float *gn_part, *mstat, *wg_part, *misc;
    size_t stat_part_floats = 0;          // capacity of stat_part
    int* state;                                    // [0] = 1 after a train-mode forward
    size_t bytes;

    Workspace(void* base, int N, int H, int W, int f, bool train);
    Act<T> make_act(Arena& ar, int lvl, int C);
};

}  // namespace mrisr
