set -x
CMD="python bench.py --steps 1 --warmup 1 --frames 48 --no-cpu-baseline"
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_tsdf.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_tsdf.log
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_b.json 2> gpurun_out/bench_b.err
B2R_TSDF_BATCH=16 timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_b16.json 2> gpurun_out/bench_b16.err
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tsdf_fuse -s 10 -c 2 -o gpurun_out/prof_fuse2 $CMD > gpurun_out/ncu2.log 2>&1 ; \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tsdf_pack -s 10 -c 2 -o gpurun_out/prof_pack2 $CMD > gpurun_out/ncu3.log 2>&1
