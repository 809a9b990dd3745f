// Host replay of the incremental unit addressing of gn_bwd_fused_kernel (csrc/gn_bwd.cu, -DMRISR_GNBWD_INCR builds).
// Walks every (cluster rank, thread) of one unit through the kernel's issue / reload / apply order with the tracker of
// csrc/gn_bwd_trk.h and compares each operand's padded-pixel index with the direct Act::pix form the default build uses.
// Also checks that phase 1 issues and phase 2 applies every unit of the thread exactly once.
// Usage: gn_bwd_trk_replay  (exit code 0 = all configurations agree; prints the first mismatch otherwise)
#include <cstdio>
#include <algorithm>
#include <cstdlib>
#include <vector>
#include "gn_bwd_trk.h"

using namespace mrisr;

static long pix(int n, int H, int W, int y, int x) { return (long)n * (H + 2) * (W + 2) + (long)(y + 1) * (W + 2) + (x + 1); }

struct Cfg { int H, W, P, FT, ncv, cs, K, S; };

static int fails = 0;
static void expect(bool ok, const Cfg& c, const char* what, int rank, int tid, int i) {
    if (ok) return;
    if (fails++ < 5)
        fprintf(stderr, "mismatch (%s): H=%d W=%d P=%d FT=%d ncv=%d cs=%d K=%d S=%d rank=%d tid=%d unit=%d\n", what, c.H, c.W, c.P,
                c.FT, c.ncv, c.cs, c.K, c.S, rank, tid, i);
}

static void replay(const Cfg& c) {
    const int n = 3;                                    // any sample: bases and direct indices carry the same term
    const int PM = c.P == 4 ? 2 : 1;
    const int UW = c.W / PM, UH = c.H / PM;
    const int rows_pc = (UH + c.cs - 1) / c.cs;
    const int pstep = c.FT / c.ncv;
    const int pitch = c.W + 2, hpitch = UW + 2, gpitch = 2 * c.W + 2;
    const TrkStep ts = trk_steps(pstep, UW, pitch, gpitch);
    const int K = c.K, S = c.S;
    for (int rank = 0; rank < c.cs; ++rank) {
        const int r0 = std::min(UH, rank * rows_pc), r1 = std::min(UH, r0 + rows_pc);
        const int npx = (r1 - r0) * UW;
        const long rbase = pix(n, c.H, c.W, PM * r0, 0);
        const long hbase = pix(n, UH, UW, r0, 0);
        const long gbase = pix(n, 2 * c.H, 2 * c.W, 2 * r0, 0);
        for (int tid = 0; tid < c.FT; ++tid) {
            const int p0 = tid / c.ncv;
            const int nu = p0 < npx ? (npx - p0 + pstep - 1) / pstep : 0;
            std::vector<int> issued(nu, 0), reissued(nu, 0), applied(nu, 0);
            auto check = [&](int i, const UnitTrk& t, const char* what) {
                const int pi = p0 + i * pstep;
                const int uy = pi / UW + r0, ux = pi % UW;
                for (int q = 0; q < c.P; ++q) {
                    const int y = PM * uy + (q >> 1), x = PM * ux + (q & 1);
                    expect(pix(n, c.H, c.W, y, x) == rbase + PM * t.px + (q >> 1) * pitch + (q & 1), c, what, rank, tid, i);
                }
                expect(t.ux == ux, c, "column", rank, tid, i);
                if (c.P == 4) expect(pix(n, UH, UW, uy, ux) == hbase + t.hpx, c, "pooled", rank, tid, i);
                if (c.P == 1)
                    for (int ij = 0; ij < 4; ++ij)
                        expect(pix(n, 2 * c.H, 2 * c.W, 2 * uy + (ij >> 1), 2 * ux + (ij & 1)) ==
                                   gbase + 2 * t.gpx + (ij >> 1) * gpitch + (ij & 1), c, "pixel-shuffle", rank, tid, i);
            };
            // slot handles of the incremental build, in slot units: advance by one, wrap over the S streaming slots
            auto slot_of = [&](int i) { return i < K ? i : K + (i - K) % S; };
            auto hnext = [&](int h) { ++h; return h == K + S ? K : h; };
            // ---- phase 1 (same order as the kernel)
            UnitTrk ti = trk_at(p0, UW, pitch, hpitch, gpitch);
            int hi = 0, hf = 0;
            for (int i = 0; i < S; ++i)
                if (i < nu) {
                    check(i, ti, "issue"); issued[i]++;
                    expect(hi == slot_of(i), c, "issue slot", rank, tid, i);
                    trk_fwd(ti, ts, UW); hi = hnext(hi);
                }
            for (int i = 0; i < nu; ++i) {
                expect(hf == slot_of(i), c, "fetch slot", rank, tid, i);
                hf = hnext(hf);
                if (i + S < nu) {
                    check(i + S, ti, "issue"); issued[i + S]++;
                    expect(hi == slot_of(i + S), c, "issue slot", rank, tid, i + S);
                    trk_fwd(ti, ts, UW); hi = hnext(hi);
                }
            }
            const int nstream = nu > K ? nu - K : 0;
            UnitTrk ta = ti;
            for (int t = 0; t < S; ++t)
                if (t < nstream) { trk_bwd(ti, ts, UW); check(nu - 1 - t, ti, "reload"); reissued[nu - 1 - t]++; }
            // ---- phase 2
            UnitTrk tc = trk_at(p0, UW, pitch, hpitch, gpitch);
            hf = 0;
            for (int i = 0; i < K && i < nu; ++i) {
                check(i, tc, "apply cached"); applied[i]++;
                expect(hf == i, c, "cached slot", rank, tid, i);
                trk_fwd(tc, ts, UW); ++hf;
            }
            hf = K;
            for (int t = 0; t < nstream; ++t) {
                expect(hf == K + t % S, c, "streamed slot", rank, tid, nu - 1 - t);
                hf = hnext(hf);
                trk_bwd(ta, ts, UW);
                check(nu - 1 - t, ta, "apply streamed");
                applied[nu - 1 - t]++;
                if (t + S < nstream) { trk_bwd(ti, ts, UW); check(nu - 1 - (t + S), ti, "reload"); reissued[nu - 1 - (t + S)]++; }
            }
            for (int i = 0; i < nu; ++i) {
                expect(issued[i] == 1, c, "issued once", rank, tid, i);
                expect(applied[i] == 1, c, "applied once", rank, tid, i);
                expect(reissued[i] == (i >= K ? 1 : 0), c, "reloaded once", rank, tid, i);
            }
        }
    }
}

int main() {
    const Cfg cfgs[] = {
        // planar planes of the U-Net levels (ncv = 1) at 256^2 / 128^2, two CTAs per cluster, streaming and all-resident
        {256, 256, 1, 512, 1, 2, 11, 4}, {256, 256, 1, 512, 1, 2, 64, 4}, {256, 256, 4, 512, 1, 2, 3, 2}, {128, 128, 1, 512, 1, 2, 16, 4},
        // NHWC samples (ncv = C / 8): pstep below and above a row, odd sizes, one-row tensors, clusters of 1 / 4
        {64, 64, 1, 512, 16, 2, 5, 4}, {32, 32, 1, 512, 32, 2, 0, 4}, {16, 16, 1, 256, 64, 2, 2, 8}, {16, 16, 4, 512, 64, 1, 1, 2},
        {30, 22, 1, 256, 2, 4, 1, 4}, {10, 8, 1, 512, 1, 2, 0, 4}, {2, 2, 1, 256, 4, 1, 0, 4}, {6, 18, 4, 256, 4, 2, 0, 2},
        {512, 512, 1, 512, 1, 2, 11, 4}, {320, 320, 4, 512, 1, 2, 3, 2}, {48, 40, 1, 512, 8, 2, 3, 8}, {5, 7, 1, 256, 1, 2, 1, 4},
    };
    for (const Cfg& c : cfgs) replay(c);
    if (fails) { fprintf(stderr, "%d mismatches\n", fails); return 1; }
    printf("ok %zu configurations\n", sizeof(cfgs) / sizeof(cfgs[0]));
    return 0;
}
