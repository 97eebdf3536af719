set -x
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_dyn.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_dyn.log
timeout 900 python tools/shard_probe.py --xls 4 --worlds 1,8 --batches 32 > gpurun_out/shard_probe.json 2> gpurun_out/shard_probe.err; echo "rc=$?" >> gpurun_out/shard_probe.err
