"""Per-layer combined roofline from a `bench.py --layer-report` table:
    python tools/layer_roofline.py profiles/r01_layers_v7.txt [batch] > profiles/...md
For every tagged conv / weight-gradient launch: algorithmic FLOPs (2*Cin*Cout*k^2*H*W*N), algorithmic HBM bytes (each operand
read once, the result written once, bf16), ideal time = max(FLOPs / sustained bf16 peak, bytes / HBM copy rate) with the peaks
of MEASURED_PEAKS.json, and the measured live CUDA-event time (the per-layer combined roofline).  For the
tagged elementwise kernels the byte model is bench.tag_bytes (minimum passes)."""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    path = sys.argv[1]
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    pk = bench.peaks()
    tf, bw = pk["bf16_tflops_sustained"] * 1e12, pk["hbm_gbs"] * 1e9
    rows, tot_ms, tot_ideal = [], 0.0, 0.0
    for line in open(path):
        m = re.match(r"(\S+)\s+calls/step\s+(\d+)\s+ms/step\s+([\d.]+)", line)
        if not m:
            continue
        tag, calls, ms = m.group(1), int(m.group(2)), float(m.group(3))
        tot_ms += ms
        fl = bench.tag_flops(tag, n)
        by = 0.0
        if fl > 0:
            mm = re.match(r"(\w+):k(\d):(\d+)>(\d+)@(\d+)x(\d+)", tag)
            ci, co, h, w = int(mm.group(3)), int(mm.group(4)), int(mm.group(5)), int(mm.group(6))
            by = 2.0 * (ci + co) * h * w * n                    # conv: read Cin, write Cout; wgrad: read Cin and Cout
        else:
            by = bench.tag_bytes(tag, n)
        if fl == 0 and by == 0:
            rows.append((tag, calls, ms, None, None, None, ""))
            continue
        t_t, t_b = fl / tf * 1e3, by / bw * 1e3                  # ms per call
        ideal = max(t_t, t_b) * calls
        tot_ideal += ideal
        rows.append((tag, calls, ms, fl * calls, by * calls, ideal, "tensor" if t_t >= t_b else "hbm"))
    print(f"# Per-layer combined roofline of `{os.path.basename(path)}` (batch {n}; peaks: {pk['bf16_tflops_sustained']:.0f} TFLOP/s sustained bf16, "
          f"{pk['hbm_gbs']:.0f} GB/s HBM copy, {pk['source']})\n")
    print("ideal = max(FLOPs / tensor peak, algorithmic bytes / HBM rate) per launch; frac = ideal / measured.  Kernels without a work "
          "model (helpers whose bytes count against the conv roofline, SURVEY.md 8(d)) have no ideal.\n")
    print("| kernel tag | calls/step | measured ms/step | GFLOP/step | MB/step | ideal ms/step | bound | frac |")
    print("|---|---|---|---|---|---|---|---|")
    for tag, calls, ms, fl, by, ideal, bound in sorted(rows, key=lambda r: -r[2]):
        if ideal is None:
            print(f"| {tag} | {calls} | {ms:.4f} | | | | | |")
        else:
            print(f"| {tag} | {calls} | {ms:.4f} | {fl / 1e9:.1f} | {by / 1e6:.0f} | {ideal:.4f} | {bound} | {ideal / ms:.2f} |")
    modelled = sum(r[2] for r in rows if r[5] is not None)
    print(f"\nSum of measured kernel time {tot_ms:.2f} ms/step, of which {modelled:.2f} ms in kernels with a work model whose ideal "
          f"sums to {tot_ideal:.2f} ms ({tot_ideal / modelled:.2f} of their measured time); the remaining {tot_ms - modelled:.2f} ms are "
          f"standalone helper passes with zero algorithmic work of their own.")


if __name__ == "__main__":
    main()
