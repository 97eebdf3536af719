set -x
for c in 4 8 16 32; do
B2R_HOST_CHUNK=$c timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_chunk$c.json 2> gpurun_out/bench_chunk$c.err; echo "rc=$?" >> gpurun_out/bench_chunk$c.err
done
timeout 600 python -m pytest tests -m gpu -x -q -k "tsdf or host or integrat" > gpurun_out/pytest_host.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_host.log
