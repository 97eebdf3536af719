set -x
CMD="python tools/bench_ops.py --only-unproject --batch 64"
timeout 300 $CMD > gpurun_out/plain_unp.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:unproject -c 60 --csv --log-file gpurun_out/launches_unp.csv $CMD > gpurun_out/ncu_u1.log 2>&1 ; \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:unproject_fill_f32 -s 2 -c 1 -o gpurun_out/prof_unp_fill $CMD > gpurun_out/ncu_u2.log 2>&1
