// Incremental unit addressing of the fused GroupNorm-backward kernel (gn_bwd.cu: the kinds with Incr<KIND>::on).
// Plain C++ so that tests/test_host_logic.py can compile it with g++ and replay the kernel's issue / apply order on the host.
//
// A thread owns the units pi = p0 + i * pstep (i = 0, 1, ...) of its CTA's rows, unit pi at (row, column) = (pi / UW, pi % UW).
// Consecutive units are d_row = pstep / UW rows and d_col = pstep % UW columns apart (one carry at most, as d_col < UW), so the
// row-pitched offsets  row * pitch + column  advance by constants:
//   px  : geometry of r / g / dr (pitch W + 2; the 2x2-window kinds address pixel 2 * px)
//   hpx : geometry of the pooled gradient (pitch UW + 2, carry 2)
//   gpx : geometry of the pixel-shuffled gradient (pitch 2W + 2; sub-pixel (0, 0) is pixel 2 * gpx)
#pragma once
#if defined(__CUDACC__)
#define MRISR_TRK_HD __host__ __device__ __forceinline__
#else
#define MRISR_TRK_HD inline
#endif

namespace mrisr {

struct TrkStep { int d_col, d_px, c_px, d_hpx, d_gpx, c_gpx; };
struct UnitTrk { int ux, px, hpx, gpx; };

inline TrkStep trk_steps(int pstep, int UW, int pitch, int gpitch) {
    const int d_row = pstep / UW, d_col = pstep % UW;
    return TrkStep{d_col, d_row * pitch + d_col, pitch - UW, d_row * (UW + 2) + d_col, d_row * gpitch + d_col, gpitch - UW};
}
MRISR_TRK_HD UnitTrk trk_at(int pi, int UW, int pitch, int hpitch, int gpitch) {
    const int row = pi / UW, col = pi - row * UW;
    return UnitTrk{col, row * pitch + col, row * hpitch + col, row * gpitch + col};
}
MRISR_TRK_HD void trk_fwd(UnitTrk& t, const TrkStep& s, int UW) {
    t.ux += s.d_col; t.px += s.d_px; t.hpx += s.d_hpx; t.gpx += s.d_gpx;
    if (t.ux >= UW) { t.ux -= UW; t.px += s.c_px; t.hpx += 2; t.gpx += s.c_gpx; }
}
MRISR_TRK_HD void trk_bwd(UnitTrk& t, const TrkStep& s, int UW) {
    t.ux -= s.d_col; t.px -= s.d_px; t.hpx -= s.d_hpx; t.gpx -= s.d_gpx;
    if (t.ux < 0) { t.ux += UW; t.px -= s.c_px; t.hpx -= 2; t.gpx -= s.c_gpx; }
}

}  // namespace mrisr
