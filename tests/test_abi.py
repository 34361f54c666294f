"""CPU-side checks of the drop-in boundary: the shared library loads, exports every symbol
include/lwc_b200.h declares, and refuses to compute without a B200 (no CPU fallback)."""
import ctypes
import json
import os
import re

import numpy as np
import pytest

import decimal_model as dm

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lwc():
    import __graft_entry__ as ge

    ge.build()
    import lwc_b200

    return lwc_b200


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "lwc_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lwc_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_boundary():
    syms = declared_symbols()
    for must in ("lwc_ctx_create", "lwc_ctx_destroy", "lwc_score_batch", "lwc_last_error_message",
                 "lwc_merge_error_codes", "lwc_stage_f64"):
        assert must in syms


def test_every_declared_symbol_is_exported(lwc):
    L = ctypes.CDLL(lwc.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(L, name), f"{name} declared in include/lwc_b200.h but not exported"


def test_no_torch_or_oracle_in_the_product_library(lwc):
    import subprocess

    needed = subprocess.run(["readelf", "-d", lwc.LIB_PATH], capture_output=True, text=True).stdout
    assert "torch" not in needed and "oracle" not in needed and "c10" not in needed


def test_merge_error_codes_host_logic(lwc):
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "tally_kats.json")))
    for case in golden["tally"]:
        if "error_codes" in case:
            assert lwc.merge_error_codes(case["error_codes"]) == case["merged_status"], case["name"]
    rng = np.random.default_rng(5)
    for _ in range(300):
        codes = [int(x) for x in rng.choice([0, 400, 401, 404, 429, 500, 502], size=rng.integers(0, 7))]
        assert lwc.merge_error_codes(codes) == dm.merge_error_codes(codes)


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="GPU present")
def test_fails_loudly_without_a_gpu(lwc):
    with pytest.raises(lwc.LwcError) as ei:
        lwc.Context(device=0)
    assert ei.value.code == 503
    assert "no CPU path" in ei.value.message


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "llm-weighted-consensus_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp", ".rs")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in text.lower() or f == "synth.py" and "oracle" not in text.lower(), os.path.join(dirpath, f)


def test_rust_sys_crate_declares_every_c_symbol():
    """rust/lwc-b200-sys cannot be compiled here (no Rust toolchain); at least keep it in step with the header."""
    rs = open(os.path.join(ROOT, "rust", "lwc-b200-sys", "src", "lib.rs")).read()
    for name in declared_symbols():
        assert re.search(r"\bfn %s\b" % name, rs), f"{name} missing from rust/lwc-b200-sys/src/lib.rs"


def test_stage_f64_is_a_host_function_and_rounds_like_a_cast(lwc):
    """lwc_stage_f64 needs no device (NULL context is fine): small arrays convert inline, large ones on threads."""
    L = lwc.load_library()
    rng = np.random.default_rng(9)
    for n in (0, 1, 1536, 5_000_001):
        src = rng.standard_normal(n) * 10.0 ** rng.integers(-20, 20, n)
        dst = np.full(n + 1, 7.0, np.float32)
        rc = L.lwc_stage_f64(None, ctypes.c_void_p(src.ctypes.data), n, ctypes.c_void_p(dst.ctypes.data))
        assert rc == 0
        assert (dst[:n] == src.astype(np.float32)).all() and dst[n] == 7.0
    assert L.lwc_stage_f64(None, None, 4, None) == 400


def test_host_staging_conversions_without_a_gpu(lwc):
    """lwc_stage_f64 / lwc_stage_f64_bf16 are host-only loops (vectorised when the CPU has AVX2): every length,
    special values and rounding ties against numpy models."""
    L = lwc.load_library()
    rng = np.random.default_rng(3)
    for n in (0, 1, 7, 8, 9, 1536, 4097):
        src = rng.standard_normal(n) * 10.0 ** rng.integers(-30, 30, n)
        if n >= 9:
            src[:9] = [np.nan, np.inf, -np.inf, 0.0, -0.0, 1e-45, 1.0 + 2.0 ** -9, 1.0 + 3 * 2.0 ** -9, 3.4e38]
        f32 = np.full(n, 7.0, np.float32)
        assert L.lwc_stage_f64(None, ctypes.c_void_p(src.ctypes.data), n, ctypes.c_void_p(f32.ctypes.data)) == 0
        with np.errstate(over="ignore"):
            want = src.astype(np.float32)
        assert np.array_equal(f32, want, equal_nan=True)
        bf = np.zeros(n, np.uint16)
        assert L.lwc_stage_f64_bf16(None, ctypes.c_void_p(src.ctypes.data), n, ctypes.c_void_p(bf.ctypes.data)) == 0
        u = want.view(np.uint32).astype(np.uint64)
        model = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)
        nan = np.isnan(want)
        assert np.array_equal(bf[~nan], model[~nan])
        back = (bf.astype(np.uint32) << 16).view(np.float32)
        assert np.all(np.isnan(back[nan]))
