set -x
CMD="python bench.py --steps 1 --warmup 1"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tsdf_pack -s 10 -c 1 -o gpurun_out/prof_pack2 $CMD > gpurun_out/ncu_k1.log 2>&1
