"""The C++ host layer (llm-weighted-consensus_b200/host/lwc_host.hpp: the reference's types for this
path + the request batcher) through its own test program, tests/cpp/test_host.cpp."""
import os
import subprocess

import pytest


@pytest.fixture(scope="module")
def exe():
    import __graft_entry__ as ge

    ge.build()
    return ge.build_host_test()


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="GPU present: covered by the gpu variant")
def test_host_layer_cpu(exe):
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "OK" in out.stdout and "503" in out.stdout


@pytest.mark.gpu
def test_host_layer_batcher_on_gpu(exe):
    out = subprocess.run([exe, "--gpu"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "OK" in out.stdout
