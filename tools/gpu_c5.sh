set -x
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo rc=$? >> gpurun_out/pytest_gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29521 tools/config5.py --ticks 4 > gpurun_out/c5_1.json 2> gpurun_out/c5_1.err; echo rc=$? >> gpurun_out/c5_1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29522 tools/config5.py --ticks 4 > gpurun_out/c5_2.json 2> gpurun_out/c5_2.err; echo rc=$? >> gpurun_out/c5_2.err
