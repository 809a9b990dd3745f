// CPU run of the staged scatter-form adjoint of bilinear x2 (csrc/upsample_bwd_slide.cuh, -DMRISR_UPSAMPLE_BWD_SLIDE builds).
// Every (sample, block, thread) of the launch grid is executed on the host, for NHWC and planar-8 tensors in fp32 and bf16,
// with and without the addend, and compared bit for bit with the gather form the shipped kernels evaluate
// (norm_act.cu: acc = sum_jy wy * (sum_jx wx * g), out = addend + acc, taps and weights from adj_weight).
// Build: nvcc -std=c++17 -I <csrc> upsample_bwd_slide_host.cu -o upsample_bwd_slide_host   (host code only; no GPU needed)
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "upsample_bwd_slide.cuh"

using namespace mrisr;

namespace mrisr {                  // referenced by common.cuh macros; unused here
void set_error(const char*, ...) {}
void count_launch() {}
}

template <typename T> static T cvt(float v);
template <> float cvt<float>(float v) { return v; }
template <> __nv_bfloat16 cvt<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
static float back(float v) { return v; }
static float back(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T>
static Act<T> make(std::vector<T>& buf, int N, int H, int W, int C, bool planar, float fill) {
    Act<T> a;
    a.N = N; a.H = H; a.W = W; a.C = C;
    const int64_t pixels = (int64_t)N * (H + 2) * (W + 2);
    if (planar) { a.ps = 8; a.cs = 8 * (pixels + 3); buf.assign((size_t)(C / 8) * a.cs, cvt<T>(fill)); }
    else { a.ps = 0; a.cs = 8; buf.assign((size_t)pixels * C, cvt<T>(fill)); }
    a.p = buf.data();
    return a;
}

// host replica of norm_act.cu: adj_weight / adj_taps (first and last destination index with a non-zero weight)
static float adj_w(int d, int i, int in_size, float scale) {
    const Lerp l = up_coord(d, in_size, scale);
    return (l.i0 == i ? 1.f - l.w1 : 0.f) + (l.i1 == i ? l.w1 : 0.f);
}
static std::vector<UpAdjTap> table(int in_size) {
    const float scale = up_scale(in_size);
    std::vector<UpAdjTap> t(in_size);
    for (int i = 0; i < in_size; ++i) {
        int lo = 0, hi = 2 * in_size - 1;
        while (lo <= hi && adj_w(lo, i, in_size, scale) == 0.f) ++lo;
        while (hi >= lo && adj_w(hi, i, in_size, scale) == 0.f) --hi;
        UpAdjTap o;
        o.lo = lo;
        o.cnt = hi - lo + 1;
        if (o.cnt > 5) { fprintf(stderr, "more than 5 taps at %d of %d\n", i, in_size); exit(2); }
        for (int j = 0; j < 5; ++j) o.w[j] = j < o.cnt ? adj_w(lo + j, i, in_size, scale) : 0.f;
        o.pad = 0.f;
        t[i] = o;
    }
    return t;
}

static int fails = 0;

template <typename T>
static void fill(Act<T>& a, unsigned seed) {
    for (int n = 0; n < a.N; ++n)
        for (int y = 0; y < a.H; ++y)
            for (int x = 0; x < a.W; ++x)
                for (int c = 0; c < a.C; ++c) {
                    seed = seed * 1664525u + 1013904223u;
                    *a.at(n, y, x, c) = cvt<T>((float)(seed >> 8) / 8388608.f - 1.f);
                }
}

template <typename T>
static void run(int N, int H, int W, int C, bool g_planar, bool o_planar, bool with_add, int rows_pb, int nthr) {
    std::vector<T> gb, ab, ob, rb;
    Act<T> g = make(gb, N, 2 * H, 2 * W, C, g_planar, 0.f);
    Act<T> add = make(ab, N, H, W, C, o_planar, 0.f);
    Act<T> out = make(ob, N, H, W, C, o_planar, -7.f);
    Act<T> ref = make(rb, N, H, W, C, o_planar, -7.f);
    fill(g, 777u + H * 31 + W);
    fill(add, 999u + C);
    if (!with_add) add.p = nullptr;
    const std::vector<UpAdjTap> ty = table(H), tx = table(W);
    for (int n = 0; n < N; ++n)
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x)
                for (int c = 0; c < C; ++c) {
                    float acc = 0.f;
                    for (int jy = 0; jy < ty[y].cnt; ++jy) {
                        float h = 0.f;
                        for (int jx = 0; jx < tx[x].cnt; ++jx)
                            h = fmaf(tx[x].w[jx], back(*g.at(n, ty[y].lo + jy, tx[x].lo + jx, c)), h);
                        acc = fmaf(ty[y].w[jy], h, acc);
                    }
                    const float a0 = with_add ? back(*add.at(n, y, x, c)) : 0.f;
                    *ref.at(n, y, x, c) = cvt<T>(a0 + acc);
                }
    const float sy = up_scale(H);
    const int blocks = (H + rows_pb - 1) / rows_pb;
    for (int n = 0; n < N; ++n)
        for (int b = 0; b < blocks; ++b)
            for (int tid = 0; tid < nthr; ++tid)
                upsample2x_bwd_slide_body<T>(g, add, out, ty.data(), tx.data(), sy, rows_pb, n, b, tid, nthr);
    size_t bad = 0;
    for (size_t i = 0; i < ob.size(); ++i) bad += memcmp(&ob[i], &rb[i], sizeof(T)) != 0;
    if (bad) {
        ++fails;
        fprintf(stderr, "mismatch: %zu of %zu elements (N=%d H=%d W=%d C=%d g_planar=%d o_planar=%d add=%d rows=%d %s)\n", bad,
                ob.size(), N, H, W, C, (int)g_planar, (int)o_planar, (int)with_add, rows_pb, sizeof(T) == 2 ? "bf16" : "fp32");
    }
}

int main() {
    const int NT = 256;
    run<__nv_bfloat16>(2, 16, 16, 128, false, false, false, 8, NT);
    run<__nv_bfloat16>(2, 32, 32, 64, false, false, false, 16, NT);
    run<__nv_bfloat16>(2, 32, 32, 32, true, false, false, 16, NT);
    run<__nv_bfloat16>(2, 64, 64, 32, true, true, true, 16, NT);
    run<__nv_bfloat16>(1, 40, 24, 16, false, true, true, 16, NT);
    run<__nv_bfloat16>(1, 7, 5, 8, false, false, true, 8, NT);
    run<__nv_bfloat16>(3, 1, 9, 24, false, false, false, 8, NT);
    run<__nv_bfloat16>(1, 9, 1, 16, false, false, true, 16, NT);
    run<__nv_bfloat16>(1, 160, 160, 16, true, true, true, 16, NT);
    run<__nv_bfloat16>(1, 33, 17, 32, true, true, false, 5, NT);
    run<float>(2, 16, 16, 64, false, false, true, 8, NT);
    run<float>(1, 13, 11, 32, false, false, false, 16, NT);
    run<float>(1, 1, 1, 8, false, false, true, 8, NT);
    run<float>(1, 2, 2, 8, false, false, false, 1, NT);
    if (fails) { fprintf(stderr, "%d configurations differ\n", fails); return 1; }
    printf("ok 14 configurations\n");
    return 0;
}
