"""One small invocation of every kernel, for `compute-sanitizer --tool memcheck|racecheck python scripts/sanitize_small.py`
(one tool per gpurun call).  Prints SANITIZE_RUN_OK when every call returned and matched the oracle loosely."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llm-weighted-consensus_b200")):
    sys.path.insert(0, p)

import lwc_b200  # noqa: E402
import oracle  # noqa: E402
from lwc_b200 import synth  # noqa: E402


def main():
    ctx = lwc_b200.Context(device=0)
    cases = [
        ("stream16", (180, 16, 256, 4), lwc_b200.PREC_FP32, 1e-5),
        ("stream16_ragged", (40, 11, 100, 3), lwc_b200.PREC_FP32, 1e-5),
        ("stream32_tf32", (60, 32, 512, 4), lwc_b200.PREC_TF32, 4e-3),
        ("direct32_x3", (60, 27, 512, 4), lwc_b200.PREC_FP32, 1e-5),
        ("generic", (10, 50, 70, 5), lwc_b200.PREC_FP32, 1e-5),
        ("tc128_tf32", (170, 128, 256, 16), lwc_b200.PREC_TF32, 4e-3),
        ("tc128_x3", (170, 100, 512, 8), lwc_b200.PREC_FP32, 1e-5),
    ]
    for name, (R, N, D, M), prec, tol in cases:
        E, V, w, _ = synth.make_batch(R, N, D, M, seed=9, edge_cases=False)
        got = ctx.score_batch(E, V, w, precision=prec)
        want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
        top = np.abs(want["conf"]).max(axis=1, keepdims=True)
        err = float(np.max(np.abs(got["conf"] - want["conf"]) / top))
        print(f"{name}: kernel={ctx.kernel_for(R, N, D, M, precision=prec)} max err/top={err:.2e}", flush=True)
        assert err <= tol * 2
    V = synth.votes(50, 16, 4, seed=1)
    w = synth.weights(50, 4, seed=1)
    ctx.score_batch(None, V, w, similarity=lwc_b200.SIM_IDENTITY)
    rng = np.random.default_rng(0)
    E = rng.standard_normal((1024, 64)).astype(np.float32)
    Vp = rng.random((3, 1024)).astype(np.float32)
    got = ctx.pool_consensus(E, Vp, np.ones(3, np.float32))
    want = oracle.pool_f64(E, Vp, np.ones(3, np.float32), method=0)
    assert np.max(np.abs(got["conf"] - want["conf"])) <= 2e-2 * want["conf"].max()
    ctx.vote_from_logprobs(rng.normal(-2, 1, (64, 20)).astype(np.float32), rng.integers(-1, 9, (64, 20)).astype(np.int32), 9)
    ctx.close()
    print("SANITIZE_RUN_OK")


if __name__ == "__main__":
    main()
