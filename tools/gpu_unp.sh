set -x
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
U="python tools/bench_ops.py --only-unproject --batch 256"
timeout 300 $U > gpurun_out/unp.json 2> gpurun_out/unp.err; echo "rc=$?" >> gpurun_out/unp.err
timeout 300 $U > gpurun_out/plain_unp.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:unproject_fused -s 2 -c 1 -o gpurun_out/prof_unp_fused $U > gpurun_out/ncu_u2.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:unproject -c 80 --csv --log-file gpurun_out/launches_unp.csv $U > gpurun_out/ncu_u1.log 2>&1
