set -x
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
for v in default c5; do
  if [ $v = default ]; then L=""; else L="$PWD/azurekinect3dreconstruction_b200/libb2recon_$v.so"; fi
  for B in 8 16; do
    B2R_LIB=$L B2R_TSDF_BATCH=$B timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/var_${v}_B$B.json 2> gpurun_out/var_${v}_B$B.err
  done
done
bash tools/gpu_prof.sh
