"""The committed golden vectors must be exactly what tests/golden/make_golden.py produces."""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def test_golden_file_is_reproducible(tmp_path):
    committed = json.load(open(os.path.join(HERE, "golden", "tally_kats.json")))
    script = open(os.path.join(HERE, "golden", "make_golden.py")).read()
    # run the generator with its output redirected into tmp_path
    work = tmp_path / "golden"
    work.mkdir()
    (work / "make_golden.py").write_text(script.replace('sys.path.insert(0, os.path.dirname(HERE))',
                                                        f'sys.path.insert(0, {HERE!r})'))
    out = subprocess.run([sys.executable, str(work / "make_golden.py")], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    fresh = json.load(open(work / "tally_kats.json"))
    assert fresh == committed
