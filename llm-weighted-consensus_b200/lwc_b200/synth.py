"""Synthetic workload generator shared by the tests and bench.py.

Counter-based PRNG (numpy Philox) keyed by (seed, stream); every array is a pure function of
its arguments.  Shapes mirror what the reference would hand the path:
  E [R][N][D] candidate-response embeddings (embeddings/response.rs:19-24, fp32 on device),
  V [R][M][N] vote rows -- one-hot (client.rs:1796-1798) or a probability vector on one
              contiguous leaf group of <= 20 choices (client.rs:1766-1791); absent voters
              (client.rs:410) are NaN rows,
  w [R][M]    voter weights, two decimals in [0.50, 2.00] (llm/mod.rs:650-660 requires > 0).
"""
import numpy as np

CONFIGS = {
    # name: (R, N, D, M)  -- BASELINE.json configs 1..4
    "cfg1": (1, 16, 1536, 4),
    "cfg2": (4096, 16, 1536, 4),
    "cfg3": (65536, 32, 3072, 4),
    "cfg4": (8192, 128, 3072, 16),
    "cfg2x4": (16384, 16, 1536, 4),  # cfg2's shape at 4x the batch: separates per-launch overhead from streaming rate
}


def _rng(seed, stream):
    return np.random.Generator(np.random.Philox(key=[int(seed) & 0xFFFFFFFFFFFFFFFF, int(stream)]))


def embeddings(R, N, D, seed=0x5EED0000, out=None, block=256):
    """Clustered embeddings: K = ceil(N/4) unit centroids per request, row i = centroid[i % K]
    + 0.35 * noise / sqrt(D), each row scaled by lognormal(0, 0.5) so normalisation matters."""
    E = out if out is not None else np.empty((R, N, D), np.float32)
    K = -(-N // 4)
    idx = np.arange(N) % K
    for r0 in range(0, R, block):
        r1 = min(R, r0 + block)
        g = _rng(seed, 1_000_000 + r0)
        c = g.standard_normal((r1 - r0, K, D), dtype=np.float32)
        c /= np.linalg.norm(c, axis=2, keepdims=True)
        x = g.standard_normal((r1 - r0, N, D), dtype=np.float32)
        x *= np.float32(0.35 / np.sqrt(D))
        x += c[:, idx, :]
        x *= g.lognormal(0.0, 0.5, (r1 - r0, N, 1)).astype(np.float32)
        E[r0:r1] = x
    return E


def votes(R, N, M, seed=0x5EED0000, absent=0.02):
    g = _rng(seed, 2)
    V = np.zeros((R, M, N), np.float32)
    width = min(N, 20)
    one_hot = g.random((R, M)) < 0.5
    pick = g.integers(0, N, (R, M))
    start = g.integers(0, N - width + 1, (R, M))
    logits = g.normal(0.0, 2.0, (R, M, width))
    p = np.exp(logits - logits.max(axis=2, keepdims=True))
    p /= p.sum(axis=2, keepdims=True)
    rr, mm = np.nonzero(~one_hot)
    cols = start[rr, mm][:, None] + np.arange(width)[None, :]
    V[rr[:, None], mm[:, None], cols] = p[rr, mm].astype(np.float32)
    rr, mm = np.nonzero(one_hot)
    V[rr, mm, pick[rr, mm]] = 1.0
    gone = g.random((R, M)) < absent
    V[gone] = np.nan
    return V


def weights(R, M, seed=0x5EED0000):
    g = _rng(seed, 3)
    return (g.integers(50, 201, (R, M)).astype(np.float32) / np.float32(100.0)).astype(np.float32)


def inject_edge_cases(E, V, w, seed=0x5EED0000, fraction=0.01):
    """~1% of requests carry one edge case: duplicate rows, a zero row, an exact tie, or no
    votes at all.  Returns {kind: [request indices]}."""
    R, N = V.shape[0], V.shape[2]
    g = _rng(seed, 4)
    n_edge = int(round(R * fraction))
    chosen = g.choice(R, size=n_edge, replace=False) if n_edge else np.array([], int)
    kinds = {"duplicate": [], "zero_row": [], "tie": [], "all_absent": []}
    names = list(kinds)
    for k, r in enumerate(chosen):
        kind = names[k % 4]
        kinds[kind].append(int(r))
        if kind == "duplicate" and E is not None and N >= 2:
            E[r, 1] = E[r, 0]
        elif kind == "zero_row" and E is not None:
            E[r, N - 1] = 0.0
        elif kind == "tie" and N >= 2:
            # two voters, equal weight, one-hot on the two copies of a duplicated row
            V[r] = np.nan
            V[r, 0] = 0.0
            V[r, 0, 0] = 1.0
            if V.shape[1] > 1:
                V[r, 1] = 0.0
                V[r, 1, 1] = 1.0
                w[r, 1] = w[r, 0]
            if E is not None:
                E[r, 1] = E[r, 0]
        elif kind == "all_absent":
            V[r] = np.nan
    return kinds


def make_batch(R, N, D, M, seed=0x5EED0000, edge_cases=True, E_out=None):
    E = embeddings(R, N, D, seed, out=E_out)
    V = votes(R, N, M, seed)
    w = weights(R, M, seed)
    kinds = inject_edge_cases(E, V, w, seed) if edge_cases else {}
    return E, V, w, kinds


def algorithmic_bytes_per_request(N, D, M):
    """4ND + 4MN + 4M (in) + 4N conf + 4N cw + 4N rank + 4M agree + 4 winner."""
    return 4 * N * D + 4 * M * N + 4 * M + 12 * N + 4 * M + 4


def algorithmic_flops_per_request(N, D, M):
    return 2 * N * N * D + 3 * N * D + 2 * M * N + 2 * N * N
