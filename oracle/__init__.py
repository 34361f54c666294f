"""ctypes loader for the CPU oracle (oracle/liblwc_oracle.so).

TEST INFRASTRUCTURE ONLY.  Importable from tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py -- never from the product package.
PARITY UNPINNED (see oracle/lwc_oracle.h): the reference has no tests or golden vectors for
this path and cannot be built here, so this restatement is pinned only against
tests/decimal_model.py and the hand-derived known answers under tests/golden/.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liblwc_oracle.so")
_SOURCES = ["lwc_oracle.c", "dec96.c", "lwc_oracle.h", "dec96.h", "Makefile"]

SIM_IDENTITY = 0
SIM_COSINE = 1


def build(force: bool = False) -> str:
    """(Re)build the oracle with gcc when a source is newer than the .so."""
    newest = max(os.path.getmtime(os.path.join(_HERE, s)) for s in _SOURCES)
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < newest:
        subprocess.check_call(["make", "-s", "-C", _HERE], env={**os.environ, "CC": "gcc"})
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = ctypes.CDLL(_SO)
        c_f = ctypes.POINTER(ctypes.c_float)
        c_d = ctypes.POINTER(ctypes.c_double)
        c_i32 = ctypes.POINTER(ctypes.c_int32)
        L.lwc_oracle_dec_op.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_size_t]
        L.lwc_oracle_dec_to_double.argtypes = [ctypes.c_char_p]
        L.lwc_oracle_dec_to_double.restype = ctypes.c_double
        L.lwc_oracle_dec_from_f32.argtypes = [ctypes.c_float, ctypes.c_char_p, ctypes.c_size_t]
        L.lwc_oracle_tally_decimal.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_char_p, ctypes.c_char_p,
                                               ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_size_t]
        L.lwc_oracle_merge_error_codes.argtypes = [ctypes.POINTER(ctypes.c_uint16), ctypes.c_int]
        score_args = [ctypes.c_int] * 5 + [c_f, c_f, c_f, ctypes.c_int, c_d, c_d, c_d, c_i32, c_i32, ctypes.c_int]
        L.lwc_oracle_score_f64.argtypes = score_args
        L.lwc_oracle_score_port.argtypes = score_args
        L.lwc_oracle_pool_f64.argtypes = [ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int, c_f, c_f, c_f,
                                          c_d, c_d, c_d, ctypes.POINTER(ctypes.c_int64), ctypes.c_int]
        L.lwc_oracle_vote_from_logprobs.argtypes = [ctypes.c_int, ctypes.c_int, c_d, c_i32, c_d]
        L.lwc_oracle_port_last_seconds.restype = ctypes.c_double
        _lib = L
    return _lib


_OPS = {"+": 0, "-": 1, "*": 2, "/": 3, "cmp": 4}


def dec_op(a: str, b: str, op: str) -> str:
    out = ctypes.create_string_buffer(128)
    rc = lib().lwc_oracle_dec_op(a.encode(), b.encode(), _OPS[op], out, len(out))
    if rc == 1:
        raise OverflowError("decimal overflow")
    if rc == 2:
        raise ZeroDivisionError("decimal division by zero")
    if rc:
        raise ValueError(f"decimal parse error ({a!r}, {b!r})")
    return out.value.decode()


def dec_to_double(a: str) -> float:
    return lib().lwc_oracle_dec_to_double(a.encode())


def dec_from_f32(x: float) -> str:
    out = ctypes.create_string_buffer(128)
    rc = lib().lwc_oracle_dec_from_f32(ctypes.c_float(x), out, len(out))
    if rc:
        raise ValueError("not finite")
    return out.value.decode()


def tally_decimal(n_choices, votes, weights):
    """votes: list (per voter) of list-of-decimal-strings or None; weights: decimal string or None.
    Returns (choice_weight, confidence, voter_confidence) as lists of strings (None where absent)."""
    M = len(votes)
    votes_csv = ";".join("-" if v is None else ",".join(v) for v in votes)
    weights_csv = ",".join("-" if w is None else w for w in weights)
    size = 64 * (n_choices + M + 2)
    bufs = [ctypes.create_string_buffer(size) for _ in range(3)]
    rc = lib().lwc_oracle_tally_decimal(n_choices, M, votes_csv.encode(), weights_csv.encode(),
                                        bufs[0], bufs[1], bufs[2], size)
    if rc:
        raise ArithmeticError(f"oracle tally failed rc={rc}")
    cw, conf, vconf = (b.value.decode().split(",") if b.value else [] for b in bufs)
    return cw, conf, [None if x == "-" else x for x in vconf]


def merge_error_codes(codes) -> int:
    arr = (ctypes.c_uint16 * max(len(codes), 1))(*codes)
    return lib().lwc_oracle_merge_error_codes(arr, len(codes))


def _ptr(a, ty):
    return a.ctypes.data_as(ctypes.POINTER(ty)) if a is not None else None


def _score(fn, E, V, w, similarity, nthreads):
    E = np.ascontiguousarray(E, dtype=np.float32)
    V = np.ascontiguousarray(V, dtype=np.float32)
    w = np.ascontiguousarray(w, dtype=np.float32)
    R, M, N = V.shape
    D = E.shape[2] if E.ndim == 3 else 0
    out = {
        "conf": np.empty((R, N), np.float64),
        "cw": np.empty((R, N), np.float64),
        "agree": np.empty((R, M), np.float64),
        "winner": np.empty((R,), np.int32),
        "rank": np.empty((R, N), np.int32),
    }
    rc = fn(R, N, D, M, similarity, _ptr(E, ctypes.c_float), _ptr(V, ctypes.c_float), _ptr(w, ctypes.c_float),
            1 if w.ndim == 2 else 0, _ptr(out["conf"], ctypes.c_double), _ptr(out["cw"], ctypes.c_double),
            _ptr(out["agree"], ctypes.c_double), _ptr(out["winner"], ctypes.c_int32),
            _ptr(out["rank"], ctypes.c_int32), nthreads)
    if rc:
        raise ArithmeticError(f"oracle score failed rc={rc}")
    return out


def score_f64(E, V, w, similarity=SIM_COSINE, nthreads=0):
    """E [R,N,D] f32, V [R,M,N] f32 (NaN first element = absent voter), w [R,M] or [M]."""
    return _score(lib().lwc_oracle_score_f64, E, V, w, similarity, nthreads)


def score_port(E, V, w, similarity=SIM_COSINE, nthreads=0):
    """Reference-typed port (Decimal tally + f64 embeddings). Adds 'seconds' = compute phase."""
    out = _score(lib().lwc_oracle_score_port, E, V, w, similarity, nthreads)
    out["seconds"] = lib().lwc_oracle_port_last_seconds()
    return out


def pool_f64(E, V, w, method=1, nthreads=0):
    E = np.ascontiguousarray(E, dtype=np.float32)
    V = np.ascontiguousarray(V, dtype=np.float32)
    w = np.ascontiguousarray(w, dtype=np.float32)
    N, D = E.shape
    M = V.shape[0]
    conf = np.empty((N,), np.float64)
    cw = np.empty((N,), np.float64)
    agree = np.empty((M,), np.float64)
    winner = ctypes.c_int64(-1)
    rc = lib().lwc_oracle_pool_f64(N, D, M, method, _ptr(E, ctypes.c_float), _ptr(V, ctypes.c_float),
                                   _ptr(w, ctypes.c_float), _ptr(conf, ctypes.c_double), _ptr(cw, ctypes.c_double),
                                   _ptr(agree, ctypes.c_double), ctypes.byref(winner), nthreads)
    if rc:
        raise ArithmeticError(f"oracle pool failed rc={rc}")
    return {"conf": conf, "cw": cw, "agree": agree, "winner": int(winner.value)}


def vote_from_logprobs(n_choices, logprobs, leaves):
    lp = np.ascontiguousarray(logprobs, dtype=np.float64)
    lf = np.ascontiguousarray(leaves, dtype=np.int32)
    vote = np.empty((n_choices,), np.float64)
    rc = lib().lwc_oracle_vote_from_logprobs(n_choices, len(lp), _ptr(lp, ctypes.c_double), _ptr(lf, ctypes.c_int32),
                                             _ptr(vote, ctypes.c_double))
    if rc:
        raise ArithmeticError("probability sum is zero")
    return vote


def table_weights(T, Q, q, top, base, minw, maxw):
    T = np.ascontiguousarray(T, dtype=np.float32)
    Q = np.ascontiguousarray(Q, dtype=np.float32)
    q = np.ascontiguousarray(q, dtype=np.float32)
    base, minw, maxw = (np.ascontiguousarray(a, dtype=np.float32) for a in (base, minw, maxw))
    rows, D = T.shape
    M = Q.shape[1]
    B = q.shape[0]
    weights = np.empty((B, M), np.float64)
    idx = np.empty((B, top), np.int32)
    sim = np.empty((B, top), np.float64)
    f = ctypes.c_float
    rc = lib().lwc_oracle_table_weights(ctypes.c_int64(rows), D, M, _ptr(T, f), _ptr(Q, f), B, _ptr(q, f), top,
                                        _ptr(base, f), _ptr(minw, f), _ptr(maxw, f), _ptr(weights, ctypes.c_double),
                                        _ptr(idx, ctypes.c_int32), _ptr(sim, ctypes.c_double))
    if rc:
        raise ValueError("invalid table limits")
    return weights, idx, sim


def num_threads() -> int:
    return lib().lwc_oracle_num_threads()
