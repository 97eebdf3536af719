# ncu evidence for profiles/: launch list + full captures of the two TSDF kernels (one ncu family per call)
set -x
CMD="python bench.py --steps 1 --warmup 1 --frames 64 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:tsdf -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1 ; \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tsdf_fuse -s 10 -c 2 -o gpurun_out/prof_fuse $CMD > gpurun_out/ncu2.log 2>&1 ; \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tsdf_pack -s 10 -c 2 -o gpurun_out/prof_pack $CMD > gpurun_out/ncu3.log 2>&1
