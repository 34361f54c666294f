"""ctypes binding of the C ABI in include/lwc_b200.h (liblwc_b200.so).

This is a thin test/bench harness over the drop-in boundary; the product is the shared
library.  There is no CPU fallback: if the library is missing, or no B200 is present,
calls raise.
"""
import ctypes
import os

import numpy as np

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("LWC_B200_LIB") or os.path.join(_PKG, "liblwc_b200.so")  # override: A/B builds of the same ABI

SIM_IDENTITY, SIM_COSINE = 0, 1
PREC_FP32, PREC_TF32X3, PREC_TF32, PREC_BF16 = 0, 1, 2, 3
MEM_HOST, MEM_DEVICE = 0, 1
ELEM_F32, ELEM_F64, ELEM_BF16 = 0, 1, 2
MAX_CHOICES, MAX_VOTERS = 128, 128
CTX_INPUTS_RESIDENT = 1


class LwcError(RuntimeError):
    """Mirrors error::ResponseError { code, message } (src/error.rs:8-13)."""

    def __init__(self, code, message):
        super().__init__(f"[{code}] {message}")
        self.code = code
        self.message = message


class BatchDesc(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in
                ("R", "N", "D", "M", "similarity", "precision", "memory", "w_per_request")]


class PoolDesc(ctypes.Structure):
    _fields_ = [("N", ctypes.c_int64), ("D", ctypes.c_int32), ("M", ctypes.c_int32), ("precision", ctypes.c_int32),
                ("memory", ctypes.c_int32)]


class Config(ctypes.Structure):
    _fields_ = [("device", ctypes.c_int32), ("chunk_requests", ctypes.c_int32), ("flags", ctypes.c_uint64),
                ("reserved", ctypes.c_uint64 * 3)]


_lib = None


def load_library():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python __graft_entry__.py` (nvcc) first; "
                          "there is no fallback implementation")
    L = ctypes.CDLL(LIB_PATH)
    vp, i32, u64 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_uint64
    L.lwc_ctx_create.argtypes = [ctypes.POINTER(Config), ctypes.POINTER(vp)]
    L.lwc_ctx_create.restype = i32
    L.lwc_ctx_destroy.argtypes = [vp]
    L.lwc_ctx_destroy.restype = None
    L.lwc_last_error_message.argtypes = [vp]
    L.lwc_last_error_message.restype = ctypes.c_char_p
    L.lwc_version.restype = ctypes.c_char_p
    for name in ("lwc_host_alloc", "lwc_device_alloc"):
        getattr(L, name).argtypes = [vp, ctypes.c_size_t, ctypes.POINTER(vp)]
        getattr(L, name).restype = i32
    for name in ("lwc_host_free", "lwc_device_free"):
        getattr(L, name).argtypes = [vp, vp]
        getattr(L, name).restype = i32
    L.lwc_memcpy.argtypes = [vp, vp, vp, ctypes.c_size_t, i32]
    L.lwc_memcpy.restype = i32
    L.lwc_ctx_stream.argtypes = [vp]
    L.lwc_ctx_stream.restype = vp
    L.lwc_stream_synchronize.argtypes = [vp, vp]
    L.lwc_stream_synchronize.restype = i32
    L.lwc_score_batch.argtypes = [vp, ctypes.POINTER(BatchDesc)] + [vp] * 9
    L.lwc_score_batch.restype = i32
    L.lwc_score_batch_typed.argtypes = [vp, ctypes.POINTER(BatchDesc), i32] + [vp] * 9
    L.lwc_score_batch_typed.restype = i32
    L.lwc_stage_f64_bf16.argtypes = [vp, vp, ctypes.c_size_t, vp]
    L.lwc_stage_f64_bf16.restype = i32
    L.lwc_score_batches.argtypes = [vp, i32, ctypes.POINTER(BatchDesc)] + [vp] * 9
    L.lwc_score_batches.restype = i32
    L.lwc_kernel_for.argtypes = [vp, ctypes.POINTER(BatchDesc)]
    L.lwc_kernel_for.restype = ctypes.c_char_p
    L.lwc_launch_count.argtypes = [vp]
    L.lwc_launch_count.restype = u64
    L.lwc_merge_error_codes.argtypes = [ctypes.POINTER(ctypes.c_uint16), i32]
    L.lwc_merge_error_codes.restype = i32
    L.lwc_vote_from_logprobs.argtypes = [vp, i32, i32, i32, vp, vp, vp, i32, vp]
    L.lwc_vote_from_logprobs.restype = i32
    L.lwc_pool_consensus.argtypes = [vp, ctypes.POINTER(PoolDesc)] + [vp] * 8
    L.lwc_pool_consensus.restype = i32
    L.lwc_pool_consensus_static.argtypes = [vp, ctypes.POINTER(PoolDesc)] + [vp] * 8
    L.lwc_pool_consensus_static.restype = i32
    L.lwc_comm_unique_id.argtypes = [vp]
    L.lwc_comm_unique_id.restype = i32
    L.lwc_comm_init.argtypes = [vp, i32, i32, vp]
    L.lwc_comm_init.restype = i32
    L.lwc_pool_p2p_export.argtypes = [vp, ctypes.c_int64, i32, vp]
    L.lwc_pool_p2p_export.restype = i32
    L.lwc_pool_p2p_attach.argtypes = [vp, vp]
    L.lwc_pool_p2p_attach.restype = i32
    L.lwc_comm_destroy.argtypes = [vp]
    L.lwc_comm_destroy.restype = i32
    L.lwc_table_load.argtypes = [vp, vp, vp, ctypes.c_int64, i32, i32]
    L.lwc_table_load.restype = i32
    L.lwc_table_weights.argtypes = [vp, vp, i32, i32, vp, vp, vp, vp, vp, vp, i32]
    L.lwc_table_weights.restype = i32
    L.lwc_table_merge.argtypes = [i32, i32, i32, i32] + [vp] * 10
    L.lwc_table_merge.restype = i32
    L.lwc_stage_f64.argtypes = [vp, vp, ctypes.c_size_t, vp]
    L.lwc_stage_f64.restype = i32
    _lib = L
    return L


def merge_error_codes(codes) -> int:
    arr = (ctypes.c_uint16 * max(len(codes), 1))(*codes)
    return load_library().lwc_merge_error_codes(arr, len(codes))


def table_merge(row_offset, idx, sim, q_rows, base, minw, maxw):
    """lwc_table_merge (host only): idx / sim [parts, prompts, top], q_rows [parts, prompts, top, M] -> weights,
    global rows and similarities of the merged `top` neighbours."""
    idx = np.ascontiguousarray(idx, dtype=np.int32)
    sim = np.ascontiguousarray(sim, dtype=np.float32)
    q_rows = np.ascontiguousarray(q_rows, dtype=np.float32)
    off = np.ascontiguousarray(row_offset, dtype=np.int64)
    base, minw, maxw = (np.ascontiguousarray(a, dtype=np.float32) for a in (base, minw, maxw))
    parts, prompts, top = idx.shape
    M = q_rows.shape[3]
    weights = np.zeros((prompts, M), np.float32)
    rows = np.zeros((prompts, top), np.int64)
    sims = np.zeros((prompts, top), np.float32)
    rc = load_library().lwc_table_merge(parts, prompts, top, M, _np_ptr(off), _np_ptr(idx), _np_ptr(sim), _np_ptr(q_rows),
                                        _np_ptr(base), _np_ptr(minw), _np_ptr(maxw), _np_ptr(weights), _np_ptr(rows),
                                        _np_ptr(sims))
    if rc:
        raise LwcError(rc, load_library().lwc_last_error_message(None).decode())
    return weights, rows, sims


def comm_unique_id() -> bytes:
    """128-byte NCCL id for lwc_comm_init; create on rank 0 and broadcast (e.g. torch.distributed)."""
    buf = ctypes.create_string_buffer(128)
    rc = load_library().lwc_comm_unique_id(buf)
    if rc:
        raise LwcError(rc, load_library().lwc_last_error_message(None).decode())
    return buf.raw


def version() -> str:
    return load_library().lwc_version().decode()


def _np_ptr(a):
    return ctypes.c_void_p(a.ctypes.data) if a is not None else None


class HostBuffer:
    """Pinned host memory from lwc_host_alloc, viewed as a numpy array."""

    def __init__(self, ctx, shape, dtype):
        self.ctx = ctx
        self.shape = tuple(int(s) for s in shape)
        self.dtype = np.dtype(dtype)
        self.nbytes = int(np.prod(self.shape, dtype=np.int64)) * self.dtype.itemsize
        p = ctypes.c_void_p()
        ctx._check(ctx.lib.lwc_host_alloc(ctx.handle, self.nbytes, ctypes.byref(p)))
        self.ptr = p.value
        raw = (ctypes.c_char * max(self.nbytes, 1)).from_address(self.ptr)
        self.array = np.frombuffer(raw, dtype=self.dtype, count=int(np.prod(self.shape, dtype=np.int64))).reshape(self.shape)

    def free(self):
        if self.ptr:
            self.array = None
            self.ctx.lib.lwc_host_free(self.ctx.handle, ctypes.c_void_p(self.ptr))
            self.ptr = None


class DeviceBuffer:
    """Device memory from lwc_device_alloc."""

    def __init__(self, ctx, nbytes):
        self.ctx = ctx
        self.nbytes = int(nbytes)
        p = ctypes.c_void_p()
        ctx._check(ctx.lib.lwc_device_alloc(ctx.handle, self.nbytes, ctypes.byref(p)))
        self.ptr = p.value

    def upload(self, arr):
        arr = np.ascontiguousarray(arr)
        assert arr.nbytes <= self.nbytes
        self.ctx._check(self.ctx.lib.lwc_memcpy(self.ctx.handle, ctypes.c_void_p(self.ptr), _np_ptr(arr), arr.nbytes, 0))
        return self

    def download(self, shape, dtype):
        out = np.empty(shape, dtype)
        assert out.nbytes <= self.nbytes
        self.ctx._check(self.ctx.lib.lwc_memcpy(self.ctx.handle, _np_ptr(out), ctypes.c_void_p(self.ptr), out.nbytes, 1))
        return out

    def free(self):
        if self.ptr:
            self.ctx.lib.lwc_device_free(self.ctx.handle, ctypes.c_void_p(self.ptr))
            self.ptr = None


class Context:
    def __init__(self, device=0, chunk_requests=0, flags=0):
        self.lib = load_library()
        cfg = Config(device=device, chunk_requests=chunk_requests, flags=flags)
        h = ctypes.c_void_p()
        rc = self.lib.lwc_ctx_create(ctypes.byref(cfg), ctypes.byref(h))
        if rc:
            raise LwcError(rc, self.lib.lwc_last_error_message(None).decode())
        self.handle = h

    def close(self):
        if self.handle:
            self.lib.lwc_ctx_destroy(self.handle)
            self.handle = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc):
        if rc:
            raise LwcError(rc, self.lib.lwc_last_error_message(self.handle).decode())

    def stream(self) -> int:
        return int(self.lib.lwc_ctx_stream(self.handle) or 0)

    def launch_count(self) -> int:
        return int(self.lib.lwc_launch_count(self.handle))

    def kernel_for(self, R, N, D, M, similarity=SIM_COSINE, precision=PREC_FP32) -> str:
        d = BatchDesc(R, N, D, M, similarity, precision, MEM_DEVICE, 1)
        return self.lib.lwc_kernel_for(self.handle, ctypes.byref(d)).decode()

    def synchronize(self, stream=None):
        self._check(self.lib.lwc_stream_synchronize(self.handle, ctypes.c_void_p(stream) if stream else None))

    def score_raw(self, desc, E, V, w, conf, cw, agree, winner, rank, stream=None):
        """Pointers (ints / c_void_p / None) straight into lwc_score_batch."""
        def vp(x):
            return x if isinstance(x, ctypes.c_void_p) or x is None else ctypes.c_void_p(x)
        self._check(self.lib.lwc_score_batch(self.handle, ctypes.byref(desc), vp(E), vp(V), vp(w), vp(conf), vp(cw),
                                             vp(agree), vp(winner), vp(rank), vp(stream)))

    def batch_list(self, descs, E, V, w, conf, cw, agree, winner, rank, stream=None):
        """Pre-builds the argument arrays of lwc_score_batches (device pointers as ints, None allowed) and returns a
        zero-argument callable that enqueues the whole list; the arrays stay alive with the callable."""
        n = len(descs)
        darr = (BatchDesc * n)(*descs)

        def parr(ptrs):
            if ptrs is None:
                return None
            return (ctypes.c_void_p * n)(*[ctypes.c_void_p(x) if x is not None else None for x in ptrs])

        arrays = [parr(a) for a in (E, V, w, conf, cw, agree, winner, rank)]
        st = ctypes.c_void_p(stream) if stream else None

        def call():
            self._check(self.lib.lwc_score_batches(self.handle, n, darr, *arrays, st))

        call.keepalive = (darr, arrays)
        return call

    def score_batch(self, E, V, w, similarity=SIM_COSINE, precision=PREC_FP32, outputs=("conf", "cw", "agree", "winner", "rank")):
        """Host-memory call: numpy in, numpy out (lwc_score_batch with LWC_MEM_HOST)."""
        V = np.ascontiguousarray(V, dtype=np.float32)
        w = np.ascontiguousarray(w, dtype=np.float32)
        R, M, N = V.shape
        if similarity == SIM_COSINE:
            E = np.ascontiguousarray(E, dtype=np.float32)
            assert E.shape[:2] == (R, N), (E.shape, V.shape)
            D = E.shape[2]
        else:
            E, D = None, 0
        desc = BatchDesc(R, N, D, M, similarity, precision, MEM_HOST, 1 if w.ndim == 2 else 0)
        out = {
            "conf": np.full((R, N), np.nan, np.float32) if "conf" in outputs else None,
            "cw": np.full((R, N), np.nan, np.float32) if "cw" in outputs else None,
            "agree": np.full((R, M), -7.0, np.float32) if "agree" in outputs else None,
            "winner": np.full((R,), -9, np.int32) if "winner" in outputs else None,
            "rank": np.full((R, N), -9, np.int32) if "rank" in outputs else None,
        }
        self.score_raw(desc, _np_ptr(E), _np_ptr(V), _np_ptr(w), _np_ptr(out["conf"]), _np_ptr(out["cw"]),
                       _np_ptr(out["agree"]), _np_ptr(out["winner"]), _np_ptr(out["rank"]))
        return {k: v for k, v in out.items() if v is not None}

    def score_batch_device(self, E, V, w, similarity=SIM_COSINE, precision=PREC_FP32):
        """Device-resident call: uploads once, runs lwc_score_batch with LWC_MEM_DEVICE, downloads."""
        V = np.ascontiguousarray(V, dtype=np.float32)
        w = np.ascontiguousarray(w, dtype=np.float32)
        R, M, N = V.shape
        D = 0
        bufs = []

        def dev(arr_or_bytes):
            b = DeviceBuffer(self, arr_or_bytes if isinstance(arr_or_bytes, int) else max(arr_or_bytes.nbytes, 4))
            bufs.append(b)
            if not isinstance(arr_or_bytes, int):
                b.upload(arr_or_bytes)
            return b

        try:
            dE = None
            if similarity == SIM_COSINE:
                E = np.ascontiguousarray(E, dtype=np.float32)
                D = E.shape[2]
                dE = dev(E)
            dV, dw = dev(V), dev(w)
            dconf, dcw = dev(4 * R * N + 4), dev(4 * R * N + 4)
            dag, dwin, drk = dev(4 * R * M + 4), dev(4 * R + 4), dev(4 * R * N + 4)
            desc = BatchDesc(R, N, D, M, similarity, precision, MEM_DEVICE, 1 if w.ndim == 2 else 0)
            self.score_raw(desc, dE.ptr if dE else None, dV.ptr, dw.ptr, dconf.ptr, dcw.ptr, dag.ptr, dwin.ptr, drk.ptr)
            self.synchronize()
            return {"conf": dconf.download((R, N), np.float32), "cw": dcw.download((R, N), np.float32),
                    "agree": dag.download((R, M), np.float32), "winner": dwin.download((R,), np.int32),
                    "rank": drk.download((R, N), np.int32)}
        finally:
            for b in bufs:
                b.free()

    def stage_f64_bf16(self, src):
        """lwc_stage_f64_bf16: float64 array -> uint16 array of bf16 bit patterns."""
        src = np.ascontiguousarray(src, dtype=np.float64)
        dst = np.empty(src.shape, np.uint16)
        self._check(self.lib.lwc_stage_f64_bf16(self.handle, _np_ptr(src), src.size, _np_ptr(dst)))
        return dst

    def score_batch_typed(self, E, e_type, V, w, precision=PREC_FP32, memory=MEM_HOST):
        """lwc_score_batch_typed: E is float64 (ELEM_F64) or uint16 bf16 bits (ELEM_BF16), host or device memory."""
        E = np.ascontiguousarray(E, dtype=np.float64 if e_type == ELEM_F64 else (np.uint16 if e_type == ELEM_BF16 else np.float32))
        V = np.ascontiguousarray(V, dtype=np.float32)
        w = np.ascontiguousarray(w, dtype=np.float32)
        R, M, N = V.shape
        D = E.shape[2]
        desc = BatchDesc(R, N, D, M, SIM_COSINE, precision, memory, 1 if w.ndim == 2 else 0)
        shapes = {"conf": ((R, N), np.float32), "cw": ((R, N), np.float32), "agree": ((R, M), np.float32),
                  "winner": ((R,), np.int32), "rank": ((R, N), np.int32)}
        vp = ctypes.c_void_p
        if memory == MEM_HOST:
            out = {k: np.zeros(sh, t) for k, (sh, t) in shapes.items()}
            self._check(self.lib.lwc_score_batch_typed(self.handle, ctypes.byref(desc), e_type, _np_ptr(E), _np_ptr(V), _np_ptr(w),
                                                       *[_np_ptr(out[k]) for k in shapes], None))
            return out
        bufs = []
        try:
            ins = [DeviceBuffer(self, max(a.nbytes, 4)).upload(a) for a in (E, V, w)]
            outs = {k: DeviceBuffer(self, int(np.prod(sh)) * 4 + 4) for k, (sh, t) in shapes.items()}
            bufs = ins + list(outs.values())
            self._check(self.lib.lwc_score_batch_typed(self.handle, ctypes.byref(desc), e_type, *[vp(b.ptr) for b in ins],
                                                       *[vp(outs[k].ptr) for k in shapes], None))
            self.synchronize()
            return {k: outs[k].download(sh, t) for k, (sh, t) in shapes.items()}
        finally:
            for b in bufs:
                b.free()

    def comm_init(self, nranks, rank, unique_id: bytes):
        buf = ctypes.create_string_buffer(unique_id, 128)
        self._check(self.lib.lwc_comm_init(self.handle, nranks, rank, buf))

    def pool_p2p_export(self, n_own, D) -> bytes:
        buf = ctypes.create_string_buffer(128)
        self._check(self.lib.lwc_pool_p2p_export(self.handle, n_own, D, buf))
        return buf.raw

    def pool_p2p_attach(self, all_handles: bytes):
        buf = ctypes.create_string_buffer(all_handles, len(all_handles))
        self._check(self.lib.lwc_pool_p2p_attach(self.handle, buf))

    def comm_destroy(self):
        self._check(self.lib.lwc_comm_destroy(self.handle))

    def pool_consensus(self, E, V, w, precision=PREC_BF16):
        """Host-memory call of lwc_pool_consensus on this rank's rows: E [N,D], V [M,N], w [M]."""
        E = np.ascontiguousarray(E, dtype=np.float32)
        V = np.ascontiguousarray(V, dtype=np.float32)
        w = np.ascontiguousarray(w, dtype=np.float32)
        N, D = E.shape
        M = V.shape[0]
        desc = PoolDesc(N, D, M, precision, MEM_HOST)
        conf = np.full((N,), np.nan, np.float32)
        cw = np.full((N,), np.nan, np.float32)
        agree = np.full((max(M, 1),), -7.0, np.float32)
        winner = ctypes.c_int64(-9)
        self._check(self.lib.lwc_pool_consensus(self.handle, ctypes.byref(desc), _np_ptr(E), _np_ptr(V), _np_ptr(w),
                                                _np_ptr(conf), _np_ptr(cw), _np_ptr(agree),
                                                ctypes.cast(ctypes.byref(winner), ctypes.c_void_p), None))
        return {"conf": conf, "cw": cw, "agree": agree[:M], "winner": int(winner.value)}

    def vote_from_logprobs(self, logprob, leaf, n_choices):
        """Host-memory call of lwc_vote_from_logprobs: [rows, top] logprobs + leaf indices -> [rows, n_choices]."""
        lp = np.ascontiguousarray(logprob, dtype=np.float32)
        lf = np.ascontiguousarray(leaf, dtype=np.int32)
        rows, top = lp.shape
        V = np.full((rows, n_choices), -7.0, np.float32)
        self._check(self.lib.lwc_vote_from_logprobs(self.handle, rows, top, n_choices, _np_ptr(lp), _np_ptr(lf),
                                                    _np_ptr(V), MEM_HOST, None))
        return V

    def table_load(self, T, Q):
        """lwc_table_load: keep a training table ([rows, D] prompt embeddings, [rows, M] per-LLM quality) on the GPU."""
        T = np.ascontiguousarray(T, dtype=np.float32)
        Q = np.ascontiguousarray(Q, dtype=np.float32)
        assert T.ndim == 2 and Q.ndim == 2 and T.shape[0] == Q.shape[0]
        self._table_shape = (T.shape[0], T.shape[1], Q.shape[1])
        self._check(self.lib.lwc_table_load(self.handle, _np_ptr(T), _np_ptr(Q), ctypes.c_int64(T.shape[0]),
                                            T.shape[1], Q.shape[1]))

    def table_weights(self, q, top, base, minw, maxw):
        """lwc_table_weights: per-prompt LLM weights from the `top` nearest rows of the loaded table."""
        q = np.ascontiguousarray(q, dtype=np.float32)
        base, minw, maxw = (np.ascontiguousarray(a, dtype=np.float32) for a in (base, minw, maxw))
        B = q.shape[0]
        M = getattr(self, "_table_shape", (0, 0, base.size))[2]
        weights = np.zeros((B, M), np.float32)
        idx = np.zeros((B, max(top, 1)), np.int32)
        sim = np.zeros((B, max(top, 1)), np.float32)
        self._check(self.lib.lwc_table_weights(self.handle, _np_ptr(q), B, top, _np_ptr(base), _np_ptr(minw),
                                               _np_ptr(maxw), _np_ptr(weights), _np_ptr(idx), _np_ptr(sim), MEM_HOST))
        return weights, idx, sim

    def table_weights_device(self, q_ptr, n_prompts, top, base, minw, maxw, weights_ptr, idx_ptr=None, sim_ptr=None):
        """lwc_table_weights with device pointers (ints): asynchronous on the context's stream."""
        base, minw, maxw = (np.ascontiguousarray(a, dtype=np.float32) for a in (base, minw, maxw))
        vp = lambda x: None if x is None else ctypes.c_void_p(x)
        self._check(self.lib.lwc_table_weights(self.handle, vp(q_ptr), n_prompts, top, _np_ptr(base), _np_ptr(minw),
                                               _np_ptr(maxw), vp(weights_ptr), vp(idx_ptr), vp(sim_ptr), MEM_DEVICE))

    def stage_f64(self, src, dst):
        src = np.ascontiguousarray(src, dtype=np.float64)
        assert dst.dtype == np.float32 and dst.size >= src.size
        self._check(self.lib.lwc_stage_f64(self.handle, _np_ptr(src), src.size, _np_ptr(dst)))
