"""Kernel-only sweep over candidate-pool sizes between the BASELINE shapes (device-resident inputs, one launch per batch,
CUDA events): where each route stands against the measured HBM peak when the pool does not fill its tile.
   usage: r2_pool_sizes.py [N,N,...] [voters] [D]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llm-weighted-consensus_b200")):
    sys.path.insert(0, p)
import torch
import bench
import lwc_b200
from lwc_b200 import synth

def main():
    torch.cuda.set_device(0)
    peak, _ = bench.load_peaks()
    D, M = int(sys.argv[3]) if len(sys.argv) > 3 else 3072, int(sys.argv[2]) if len(sys.argv) > 2 else 8
    sizes = [int(x) for x in (sys.argv[1].split(",") if len(sys.argv) > 1 else "8,16,24,32,40,48,64,80,96,128".split(","))]
    for N in sizes:
        R = max(256, int(1.6e9 / (N * D * 4)))
        for prec, pname in ((lwc_b200.PREC_FP32, "fp32"), (lwc_b200.PREC_TF32, "tf32")):
            ms, kernel, par = bench.timed_side_shape(torch, lwc_b200, synth, None, None, R, N, D, M, prec, 5, 3)
            gbs = R * synth.algorithmic_bytes_per_request(N, D, M) / (ms * 1e-3) / 1e9
            print(json.dumps({"N": N, "D": D, "M": M, "R": R, "precision": pname, "kernel": kernel, "ms": round(ms, 4), "gbs": round(gbs, 1),
                              "frac": round(gbs / peak, 3), "parity_ok": bool(par and par.get("winner_agreement", 0) == 1.0), "max_err": par and par.get("max_conf_err_over_top")}), flush=True)

if __name__ == "__main__":
    main()
