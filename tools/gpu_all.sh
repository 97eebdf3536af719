set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python tools/bench_ops.py > gpurun_out/bench_ops.json 2> gpurun_out/bench_ops.err; echo "rc=$?" >> gpurun_out/bench_ops.err
A="python tools/bench_ops.py --frames 48 --pairs 2"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"mesh|extract|sort|Sort|scan|Scan|scatter|pack_keys|normals" -c 400 --csv --log-file gpurun_out/launches_ops.csv $A > gpurun_out/ncu_o1.log 2>&1
