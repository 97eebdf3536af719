# Round-end verification: GPU tests, smoke, both bench arms, ops bench, then the ncu evidence of the bench command
set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "rc=$?" >> gpurun_out/bench.err
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "rc=$?" >> gpurun_out/bench_ref.err
timeout 600 python tools/shard_probe.py --xls 4 --batches 32 > gpurun_out/shard_probe.json 2> gpurun_out/shard_probe.err; echo "rc=$?" >> gpurun_out/shard_probe.err
bash tools/gpu_prof.sh
