# first GPU pass over the batched TSDF pipeline: tests, then a small sweep of the two tuning knobs
set -x
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
for B in 8 1 4 16; do for S in 2 3; do
  B2R_TSDF_BATCH=$B B2R_FUSE_STAGES=$S timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_B${B}_S${S}.json 2> gpurun_out/bench_B${B}_S${S}.err; echo "rc=$?" >> gpurun_out/bench_B${B}_S${S}.err
done; done
