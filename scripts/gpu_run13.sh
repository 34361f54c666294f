set -x
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest13a.log 2>&1; echo "alone rc=$?" >> gpurun_out/pytest13a.log; tail -15 gpurun_out/pytest13a.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest13b.log 2>&1; echo "all rc=$?" >> gpurun_out/pytest13b.log; tail -8 gpurun_out/pytest13b.log
