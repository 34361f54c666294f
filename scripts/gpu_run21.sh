set -x
for c in 0 128 340 1024 2048 4096; do
LWC_CHUNK_REQUESTS=$c timeout 150 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/e2e_$c.log 2>&1
echo "chunk=$c: $(grep -o '"e2e": {"value": [0-9.]*' gpurun_out/e2e_$c.log) $(grep -o '"wall_ms_per_step": [0-9.]*' gpurun_out/e2e_$c.log)"
done
nvidia-smi --query-gpu=pcie.link.gen.current,pcie.link.width.current,pcie.link.gen.max --format=csv
