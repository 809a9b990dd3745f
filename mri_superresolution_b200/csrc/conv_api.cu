// C ABI of the convolution building block (include/mrisr.h: mrisr_conv2d_*): one conv / dgrad / wgrad over padded NHWC
// bf16 tensors with either engine.  Used by the parity tests to check the tcgen05 kernels layer by layer against the
// CUDA-core kernels and a torch fp32 reference, and by the VGG perceptual branch.
#include "../../include/mrisr.h"
#include "kernels.cuh"

using namespace mrisr;
using bf16 = __nv_bfloat16;

namespace {
struct Carve {
    char* base; size_t used = 0;
    explicit Carve(void* b) : base((char*)b) {}
    void* take(size_t bytes) { void* p = base ? base + used : nullptr; used += (bytes + 255) & ~(size_t)255; return p; }
};
struct Scratch {
    bf16 *wf, *wd; float* part; size_t bytes;
    Scratch(void* b, int N, int H, int W, int C1, int C2, int Cout, int ks) {
        Carve c(b);
        const size_t wel = (size_t)(C1 + C2) * Cout * ks * ks;
        wf = (bf16*)c.take(wel * sizeof(bf16));
        wd = (bf16*)c.take(wel * sizeof(bf16));
        size_t pe = (size_t)N * conv_generic_stat_slots(H, W, Cout) * kGroups * 2;
        pe = pe > (size_t)conv_tc_stat_elems(N, H, W) ? pe : (size_t)conv_tc_stat_elems(N, H, W);
        pe = pe > (size_t)conv_ring_stat_elems(N, H, W, C1 + C2, Cout) ? pe : (size_t)conv_ring_stat_elems(N, H, W, C1 + C2, Cout);
        size_t we = wgrad_generic_scratch_elems(N, H, W, C1 + C2, Cout, ks);
        if (wgrad_tc_supported(H, W, C1, C2, Cout, ks)) {
            const size_t t = wgrad_tc_scratch_elems(N, H, W, C1, C2, Cout, ks);
            we = we > t ? we : t;
        }
        if (wgrad_ring_supported(H, W, C1, C2, Cout, ks)) {
            const size_t t = wgrad_ring_scratch_elems(N, H, W, C1 + C2, Cout, ks);
            we = we > t ? we : t;
        }
        if (wgrad_rows_supported(N, H, W, C1, C2, Cout, ks)) {
            const size_t t = wgrad_rows_scratch_elems(N, H, W, C1 + C2, Cout);
            we = we > t ? we : t;
        }
        part = (float*)c.take((pe > we ? pe : we) * sizeof(float));
        bytes = c.used;
    }
};
// engine 0 / 1: `p` is the padded origin of an NHWC tensor.  engine 2: `p` is the start of a planar-8 allocation
// (C/8 planes of planar_plane() pixels x 8 channels, pixel 0 sits planar_guard() pixels into each plane).
Act<bf16> act(const void* p, int N, int H, int W, int C, int engine = 0) {
    Act<bf16> a{(bf16*)p, N, H, W, C};
    if (engine == 2 || engine == 3) {
        a.ps = 8;
        a.cs = planar_plane(a.rows(), W) * 8;
        a.p = (bf16*)p + planar_guard(W) * 8;
    }
    return a;
}
}  // namespace

extern "C" {

int mrisr_planar_geometry(int N, int H, int W, long long* guard_pixels, long long* plane_pixels) {
    if (N <= 0 || H <= 0 || W <= 0 || !guard_pixels || !plane_pixels) { set_error("planar_geometry: bad argument"); return MRISR_E_BADARG; }
    *guard_pixels = planar_guard(W);
    *plane_pixels = planar_plane((int64_t)N * (H + 2) * (W + 2), W);
    return 0;
}

size_t mrisr_conv2d_scratch_bytes(int N, int H, int W, int C1, int C2, int Cout, int ks) {
    if (N <= 0 || H <= 0 || W <= 0 || C1 <= 0 || C2 < 0 || Cout <= 0 || (ks != 1 && ks != 3)) return 0;
    const size_t a = Scratch(nullptr, N, H, W, C1, C2, Cout, ks).bytes;
    const size_t b = Scratch(nullptr, N, H, W, Cout, 0, C1 + C2, ks).bytes;      // dgrad direction
    return a > b ? a : b;
}

int mrisr_conv2d_bf16(const void* src1, int C1, const void* src2, int C2, const float* w, const float* bias, void* dst,
                      float* stats, void* scratch, size_t scratch_bytes, int N, int H, int W, int Cout, int ks,
                      int transposed, int engine, void* stream) {
    if (!src1 || !w || !dst || !scratch || N <= 0 || H <= 0 || W <= 0 || (ks != 1 && ks != 3)) { set_error("conv2d: bad argument"); return MRISR_E_BADARG; }
    if (transposed && (src2 || bias || stats)) { set_error("conv2d: dgrad takes one source, no bias, no statistics"); return MRISR_E_BADARG; }
    cudaStream_t s = (cudaStream_t)stream;
    const int Cin = C1 + (src2 ? C2 : 0);
    Scratch sc(scratch, N, H, W, C1, src2 ? C2 : 0, Cout, ks);
    if (sc.bytes > scratch_bytes) { set_error("conv2d: scratch too small (%zu > %zu)", sc.bytes, scratch_bytes); return MRISR_E_WORKSPACE; }
    // forward weight dims: [Cout][Cin] normally; for dgrad the caller passes the FORWARD weight [Cin_here = C1][Cout_here]
    const int wCout = transposed ? C1 : Cout, wCin = transposed ? Cout : Cin;
    ConvArgs<bf16> a;
    a.src1 = act(src1, N, H, W, C1, engine);
    a.src2 = src2 ? act(src2, N, H, W, C2, engine) : Act<bf16>{nullptr, 0, 0, 0, 0};
    a.bias = bias; a.dst = act(dst, N, H, W, Cout, engine); a.stat_part = stats ? sc.part : nullptr; a.ks = ks;
    int rc;
    if (engine == 2) {
        if ((rc = launch_pack_weights_ring(w, sc.wf, sc.wd, wCout, wCin, ks, s))) return rc;
        a.w = transposed ? sc.wd : sc.wf;
        if ((rc = launch_conv_ring(a, s))) return rc;
        if (stats) return launch_gn_finalize_ring(sc.part, stats, N, H, W, Cin, Cout, (double)(Cout / kGroups) * H * W, s);
        return 0;
    }
    if (engine == 1) {
        if ((rc = launch_pack_weights_tc(w, sc.wf, sc.wd, wCout, wCin, ks, s))) return rc;
        a.w = transposed ? sc.wd : sc.wf;
        if ((rc = launch_conv_tc(a, s))) return rc;
        if (stats) return launch_gn_finalize_tc(sc.part, stats, N, H, W, (double)(Cout / kGroups) * H * W, s);
        return 0;
    }
    if ((rc = launch_pack_weights<bf16>(w, sc.wf, sc.wd, wCout, wCin, ks, s))) return rc;
    a.w = transposed ? sc.wd : sc.wf;
    if ((rc = launch_conv_generic<bf16>(a, s))) return rc;
    if (stats) return launch_gn_finalize(sc.part, stats, N, conv_generic_stat_slots(H, W, Cout), (double)(Cout / kGroups) * H * W, s);
    return 0;
}

int mrisr_conv2d_wgrad_bf16(const void* src1, int C1, const void* src2, int C2, const void* dy, float* dw, void* scratch,
                            size_t scratch_bytes, int N, int H, int W, int Cout, int ks, int engine, void* stream) {
    if (!src1 || !dy || !dw || !scratch || N <= 0 || H <= 0 || W <= 0 || (ks != 1 && ks != 3)) { set_error("wgrad: bad argument"); return MRISR_E_BADARG; }
    cudaStream_t s = (cudaStream_t)stream;
    Scratch sc(scratch, N, H, W, C1, src2 ? C2 : 0, Cout, ks);
    if (sc.bytes > scratch_bytes) { set_error("wgrad: scratch too small (%zu > %zu)", sc.bytes, scratch_bytes); return MRISR_E_WORKSPACE; }
    WgradArgs<bf16> a;
    a.src1 = act(src1, N, H, W, C1, engine);
    a.src2 = src2 ? act(src2, N, H, W, C2, engine) : Act<bf16>{nullptr, 0, 0, 0, 0};
    a.dy = act(dy, N, H, W, Cout, engine); a.dw = dw; a.partial = sc.part; a.ks = ks;
    if (engine == 3) return launch_wgrad_rows(a, s);          // row-structured weight gradient (planar-8, as engine 2)
    if (engine == 2) return launch_wgrad_ring(a, s);
    return engine == 1 ? launch_wgrad_tc(a, s) : launch_wgrad_generic<bf16>(a, s);
}

}  // extern "C"
