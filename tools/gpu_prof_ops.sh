# ncu evidence for the non-headline rows: unprojection (a3-a6), extraction (a10/a11), odometry/ICP (a12/a13)
set -x
U="python tools/bench_ops.py --only-unproject --batch 64"
A="python tools/bench_ops.py --frames 48 --pairs 2"
timeout 300 $U > gpurun_out/plain_unp.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:unproject -c 80 --csv --log-file gpurun_out/launches_unp.csv $U > gpurun_out/ncu_u1.log 2>&1 ; \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:unproject_fill_f32 -s 2 -c 1 -o gpurun_out/prof_unp_fill $U > gpurun_out/ncu_u2.log 2>&1 ; \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:unproject_count -s 2 -c 1 -o gpurun_out/prof_unp_count $U > gpurun_out/ncu_u3.log 2>&1
timeout 300 $A > gpurun_out/plain_ops.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"mesh|extract|odo|icp|image|pyr|sort|Sort|scan|Scan" -c 1500 --csv --log-file gpurun_out/launches_ops.csv $A > gpurun_out/ncu_o1.log 2>&1 ; \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"mesh_fill|mesh_count|extract_points" -c 6 -o gpurun_out/prof_extract $A > gpurun_out/ncu_o2.log 2>&1
