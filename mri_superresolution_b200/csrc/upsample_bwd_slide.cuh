// Adjoint of bilinear x2 (align_corners=True) in scatter-along-rows form (staged build, -DMRISR_UPSAMPLE_BWD_SLIDE; the default
// is upsample2x_bwd_tiled_kernel in norm_act.cu).
//
// The tiled kernel evaluates, for every source-gradient vector, cnt_y x cnt_x (16 - 25) taps of the destination gradient:
// the horizontal reduction of a destination row is recomputed for each of the ~2 source rows it feeds.  Here a thread owns
// source columns (horizontal tap window and weights in registers) and walks the DESTINATION rows d the block's source rows
// gather from, in ascending order: one horizontal reduction h(d) per row, scattered with the forward weights (1 - w1, w1) of
// up_coord(d) into two running accumulators (source rows i0(d) and i0(d) + 1); when i0 advances the finished row is stored.
// Per source row the contributions arrive in ascending d, i.e. the same summation order and the same weights as the gather
// form (adj_weight), so the results are bit-identical to it.
// The body is a plain function of (sample, block, thread): tests/native/upsample_bwd_slide_host.cu runs it on the CPU.
#pragma once
#include "common.cuh"

namespace mrisr {

// tap window of one source index along one axis: destination indices [lo, lo + cnt) with their weights (32 bytes; same layout
// as norm_act.cu's AdjTap5, whose device tables are passed here)
struct UpAdjTap { int lo, cnt; float w[5]; float pad; };

template <typename T>
__host__ __device__ __forceinline__ void upsample2x_bwd_slide_body(const Act<T>& ddst, const Act<T>& addend, const Act<T>& dsrc,
                                                                   const UpAdjTap* tab_y, const UpAdjTap* tab_x, float sy,
                                                                   int rows_pb, int n, int block, int tid, int nthr) {
    const RowMap m = row_map(dsrc, nthr, tid);
    if (!m.active) return;
    const int y0 = block * rows_pb, y1 = y0 + rows_pb < dsrc.H ? y0 + rows_pb : dsrc.H;
    if (y0 >= y1) return;
    const int d_lo = tab_y[y0].lo, d_hi = tab_y[y1 - 1].lo + tab_y[y1 - 1].cnt;       // destination rows [d_lo, d_hi)
    const int psg = ddst.ps ? ddst.ps : ddst.C, pso = dsrc.ps ? dsrc.ps : dsrc.C;
    const int grs = (ddst.W + 2) * psg, ors = (dsrc.W + 2) * pso;                    // row strides in elements (< 2^31: host check)
    const int pad_ = addend.p ? (addend.ps ? addend.ps : addend.C) : 0;
    const int ars = addend.p ? (addend.W + 2) * pad_ : 0;
    for (int x = m.x0; x < dsrc.W; x += m.xstep) {
        const UpAdjTap tx = tab_x[x];
        const T* gp = ddst.at(n, 0, tx.lo, m.cv * 8);                                // first horizontal tap of destination row 0
        T* op = dsrc.at(n, 0, x, m.cv * 8);
        const T* ap = addend.p ? addend.at(n, 0, x, m.cv * 8) : nullptr;
        float accA[8], accB[8];
        MRISR_UNROLL
        for (int k = 0; k < 8; ++k) accA[k] = accB[k] = 0.f;
        int cur = y0;                                                                // source row accumulated in accA (accB: cur + 1)
        auto emit = [&]() {                                                          // store row `cur`, slide the accumulators
            if (cur < y1) {
                float o[8];
                if (ap) {
                    float add[8];
                    load8(ap + (int64_t)cur * ars, add);
                    MRISR_UNROLL
                    for (int k = 0; k < 8; ++k) o[k] = add[k] + accA[k];
                } else {
                    MRISR_UNROLL
                    for (int k = 0; k < 8; ++k) o[k] = 0.f + accA[k];
                }
                store8(op + (int64_t)cur * ors, o);
            }
            MRISR_UNROLL
            for (int k = 0; k < 8; ++k) { accA[k] = accB[k]; accB[k] = 0.f; }
            ++cur;
        };
        for (int d = d_lo; d < d_hi; ++d) {
            const Lerp ly = up_coord(d, dsrc.H, sy);
            while (ly.i0 > cur) emit();                                              // block-uniform; at most once per d after the first
            float h[8];
            MRISR_UNROLL
            for (int k = 0; k < 8; ++k) h[k] = 0.f;
            const T* row = gp + (int64_t)d * grs;
            MRISR_UNROLL
            for (int jx = 0; jx < 5; ++jx) {
                if (jx < tx.cnt) {
                    float v[8];
                    load8(row + jx * psg, v);
                    MRISR_UNROLL
                    for (int k = 0; k < 8; ++k) h[k] = fmaf(tx.w[jx], v[k], h[k]);
                }
            }
            // weights of row d towards its two source rows, formed as adj_weight does
            const float wa = (1.f - ly.w1) + (ly.i1 == ly.i0 ? ly.w1 : 0.f);
            const float wb = ly.i1 == ly.i0 ? 0.f : ly.w1;
            if (ly.i0 == cur) {
                MRISR_UNROLL
                for (int k = 0; k < 8; ++k) { accA[k] = fmaf(wa, h[k], accA[k]); accB[k] = fmaf(wb, h[k], accB[k]); }
            } else {                                                                 // ly.i0 == cur - 1: only its lower row is ours
                MRISR_UNROLL
                for (int k = 0; k < 8; ++k) accA[k] = fmaf(wb, h[k], accA[k]);
            }
        }
        while (cur < y1) emit();
    }
}

#ifdef __CUDACC__
template <typename T, int NT>
__global__ void __launch_bounds__(NT) upsample2x_bwd_slide_kernel(Act<T> ddst, Act<T> addend, Act<T> dsrc, const UpAdjTap* __restrict__ tab_y,
                                                                  const UpAdjTap* __restrict__ tab_x, float sy, int rows_pb) {
    upsample2x_bwd_slide_body<T>(ddst, addend, dsrc, tab_y, tab_x, sy, rows_pb, blockIdx.y, blockIdx.x, threadIdx.x, NT);
}
#endif

}  // namespace mrisr
