"""Latency of a few-request host call through lwc_score_batch(LWC_MEM_HOST), per LWC_SMALL_ZERO_COPY mode (set in the
environment before the context is made).  Prints one JSON line per (mode, R)."""
import ctypes, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "llm-weighted-consensus_b200"))

def run(mode, shapes, reps=3000):
    os.environ["LWC_SMALL_ZERO_COPY"] = str(mode)
    import lwc_b200
    from lwc_b200 import synth
    ctx = lwc_b200.Context(device=0)
    for (R, N, D, M) in shapes:
        E, V, w, _ = synth.make_batch(R, N, D, M, seed=7, edge_cases=False)
        hb = [lwc_b200.HostBuffer(ctx, a.shape, np.float32) for a in (E, V, w)]
        for b, a in zip(hb, (E, V, w)):
            b.array[...] = a
        outs = [lwc_b200.HostBuffer(ctx, sh, t) for sh, t in (((R, N), np.float32), ((R, N), np.float32), ((R, M), np.float32), ((R,), np.int32), ((R, N), np.int32))]
        desc = lwc_b200.BatchDesc(R, N, D, M, lwc_b200.SIM_COSINE, lwc_b200.PREC_FP32, lwc_b200.MEM_HOST, 1)
        vp = ctypes.c_void_p
        argv = (ctx.handle, ctypes.byref(desc), *[vp(b.ptr) for b in hb], *[vp(o.ptr) for o in outs], None)
        fn = ctx.lib.lwc_score_batch
        for _ in range(200):
            fn(*argv)
        us = np.empty(reps)
        for i in range(reps):
            t0 = time.perf_counter_ns()
            rc = fn(*argv)
            us[i] = (time.perf_counter_ns() - t0) * 1e-3
            assert rc == 0
        print(json.dumps({"mode": mode, "shape": [R, N, D, M], "median_us": round(float(np.median(us)), 2), "p10_us": round(float(np.percentile(us, 10)), 2),
                          "p99_us": round(float(np.percentile(us, 99)), 2), "conf0": float(outs[0].array[0, 0]), "winner": int(outs[3].array[0])}), flush=True)
        for b in hb + outs:
            b.free()
    ctx.close()

if __name__ == "__main__":
    shapes = [(1, 16, 1536, 4), (2, 16, 1536, 4), (1, 32, 1536, 8), (1, 4, 256, 2)]
    for mode in (0, 1, 2, 3, 4):
        run(mode, shapes)
