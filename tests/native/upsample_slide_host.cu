// CPU run of the staged sliding bilinear-x2 kernel body (csrc/upsample_slide.cuh, -DMRISR_UPSAMPLE_SLIDE builds).
// The body is a plain __host__ __device__ function of (sample, block, thread): here every (n, block, tid) of the launch grid is
// executed on the host, for NHWC and planar-8 tensors in fp32 and bf16, and compared bit for bit with the direct form the
// default kernel evaluates (norm_act.cu: upsample2x_kernel).  The zero border of the destination must stay untouched.
// Build: nvcc -std=c++17 -I <csrc> upsample_slide_host.cu -o upsample_slide_host   (host code only; no GPU needed)
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "upsample_slide.cuh"

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
    if (planar) { a.ps = 8; a.cs = 8 * (pixels + 5); buf.assign((size_t)(C / 8) * a.cs, cvt<T>(fill)); }
    else { a.ps = 0; a.cs = 8; buf.assign((size_t)pixels * C, cvt<T>(fill)); }
    a.p = buf.data();
    return a;
}

static int fails = 0;

template <typename T, int UB, bool PF = false>
static void run(int N, int H, int W, int C, bool src_planar, bool dst_planar, int rows_pb, int nthr) {
    std::vector<T> sb, db, rb;
    Act<T> src = make(sb, N, H, W, C, src_planar, 0.f);
    Act<T> dst = make(db, N, 2 * H, 2 * W, C, dst_planar, -7.f);
    Act<T> ref = make(rb, N, 2 * H, 2 * W, C, dst_planar, -7.f);
    unsigned seed = 12345u + H * 131 + W * 7 + C;
    for (int n = 0; n < N; ++n)
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x)
                for (int c = 0; c < C; ++c) {
                    seed = seed * 1664525u + 1013904223u;
                    *src.at(n, y, x, c) = cvt<T>((float)(seed >> 8) / 8388608.f - 1.f);
                }
    const float sy = up_scale(H), sx = up_scale(W);
    for (int n = 0; n < N; ++n)
        for (int Y = 0; Y < 2 * H; ++Y) {
            const Lerp ly = up_coord(Y, H, sy);
            const float h1 = ly.w1, h0 = 1.f - h1;
            for (int X = 0; X < 2 * W; ++X) {
                const Lerp lx = up_coord(X, W, sx);
                const float a1 = lx.w1, a0 = 1.f - a1;
                for (int c = 0; c < C; ++c) {
                    const float v0 = back(*src.at(n, ly.i0, lx.i0, c)), v1 = back(*src.at(n, ly.i0, lx.i1, c));
                    const float v2 = back(*src.at(n, ly.i1, lx.i0, c)), v3 = back(*src.at(n, ly.i1, lx.i1, c));
                    *ref.at(n, Y, X, c) = cvt<T>(h0 * (a0 * v0 + a1 * v1) + h1 * (a0 * v2 + a1 * v3));
                }
            }
        }
    const int blocks = (2 * H + rows_pb - 1) / rows_pb;
    for (int n = 0; n < N; ++n)
        for (int b = 0; b < blocks; ++b)
            for (int tid = 0; tid < nthr; ++tid) upsample2x_slide_body<T, UB, PF>(src, dst, sy, sx, rows_pb, n, b, tid, nthr);
    size_t bad = 0;
    for (size_t i = 0; i < db.size(); ++i) bad += memcmp(&db[i], &rb[i], sizeof(T)) != 0;
    if (bad) {
        ++fails;
        fprintf(stderr, "mismatch: %zu of %zu elements (N=%d H=%d W=%d C=%d src_planar=%d dst_planar=%d rows=%d UB=%d %s)\n", bad,
                db.size(), N, H, W, C, (int)src_planar, (int)dst_planar, rows_pb, UB * (PF ? -1 : 1), sizeof(T) == 2 ? "bf16" : "fp32");
    }
}

int main() {
    const int NT = 256;
    // the four launches of a 256^2 step, scaled down (same channel counts and layouts), odd sizes, one-row / one-column sources
    run<__nv_bfloat16, 1>(2, 16, 16, 128, false, false, 8, NT);
    run<__nv_bfloat16, 1>(2, 32, 32, 64, false, false, 16, NT);
    run<__nv_bfloat16, 1>(2, 32, 32, 32, false, true, 16, NT);
    run<__nv_bfloat16, 1>(2, 64, 64, 32, true, true, 16, NT);
    run<__nv_bfloat16, 2>(1, 64, 64, 32, true, true, 16, NT);
    run<__nv_bfloat16, 1>(1, 40, 24, 16, true, false, 16, NT);
    run<__nv_bfloat16, 1>(1, 7, 5, 8, false, false, 8, NT);
    run<__nv_bfloat16, 2>(3, 1, 9, 24, false, false, 8, NT);
    run<__nv_bfloat16, 1>(1, 9, 1, 16, false, false, 16, NT);
    run<__nv_bfloat16, 1>(1, 160, 160, 16, true, true, 16, NT);
    run<float, 1>(2, 16, 16, 64, false, false, 8, NT);
    run<float, 2>(1, 13, 11, 32, false, false, 16, NT);
    run<float, 1>(1, 1, 1, 8, false, false, 8, NT);
    // the same with the one-step-ahead prefetch of the next source row
    run<__nv_bfloat16, 1, true>(2, 32, 32, 32, false, true, 16, NT);
    run<__nv_bfloat16, 1, true>(2, 64, 64, 32, true, true, 16, NT);
    run<__nv_bfloat16, 2, true>(1, 40, 24, 16, true, false, 16, NT);
    run<__nv_bfloat16, 1, true>(1, 7, 5, 8, false, false, 8, NT);
    run<__nv_bfloat16, 1, true>(3, 1, 9, 24, false, false, 8, NT);
    run<__nv_bfloat16, 1, true>(1, 9, 1, 16, false, false, 5, NT);
    run<float, 1, true>(1, 13, 11, 32, false, false, 16, NT);
    if (fails) { fprintf(stderr, "%d configurations differ\n", fails); return 1; }
    printf("ok 20 configurations\n");
    return 0;
}
