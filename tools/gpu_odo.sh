set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
A="python tools/bench_ops.py --frames 24 --pairs 20"
timeout 600 $A > gpurun_out/ops_small.json 2> gpurun_out/ops_small.err
B2R_ODO_GRAPH=0 timeout 600 $A > gpurun_out/ops_small_nograph.json 2> gpurun_out/ops_small_nograph.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"corr_warp|reduce_kernel|icp" -c 300 --csv --log-file gpurun_out/launches_odo.csv python tools/bench_ops.py --frames 8 --pairs 3 > gpurun_out/ncu_odo.log 2>&1
