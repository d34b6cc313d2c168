"""Development tool: time mgCycle@finest and one smoothing sweep of the 3D optical-flow program on one B200 (CUDA events).
Usage: python tools/optflow3_bench.py [max_level=8]  (2^level cells per dimension)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from exastencils_b200 import OpticalFlow3DProgram
    hi = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    stream = torch.cuda.Stream()
    p = OpticalFlow3DProgram(2, hi, stream=stream.cuda_stream)
    p.init_images()
    p.init_fields()
    nx, ny, nz = p.cells()
    cells = float(nx) * ny * nz

    def timed(fn, reps):
        with torch.cuda.stream(stream):
            for _ in range(2):
                fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            stream.synchronize()
            e0.record(stream)
            for _ in range(reps):
                fn()
            e1.record(stream)
            stream.synchronize()
        return e0.elapsed_time(e1) / reps

    r0 = p.update_residual_norm()
    ms_cycle = timed(lambda: p.vcycle(), 5)
    r1 = p.update_residual_norm()
    ms_sweep = timed(lambda: p.smooth(hi, 1), 10)
    ms_res = timed(lambda: p.residual(hi), 10)
    # per cell and sweep: 6 coefficients + 3 rhs + 3 flow read + 3 flow written = 120 B; residual: 6 + 3 + 3 + 3 = 120 B
    print("cells %d^3: mgCycle %.3f ms (%.2f G unknown-cycles/s), residual norm %.3e -> %.3e after 7 cycles" %
          (nx, ms_cycle, 3.0 * cells / ms_cycle / 1e6, r0, r1))
    print("  one sweep (two colours) %.3f ms = %.0f GB/s algorithmic (%.2f of %.0f); residual %.3f ms = %.0f GB/s (%.2f)" %
          (ms_sweep, 120.0 * cells / ms_sweep / 1e6, 120.0 * cells / ms_sweep / 1e6 / peak, peak, ms_res,
           120.0 * cells / ms_res / 1e6, 120.0 * cells / ms_res / 1e6 / peak))
    p.close()


if __name__ == "__main__":
    main()
