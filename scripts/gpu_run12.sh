set -x
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest12.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest12.log; tail -25 gpurun_out/pytest12.log
