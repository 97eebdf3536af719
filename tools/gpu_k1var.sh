set -x
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_golden.py tests/test_gpu_fullsize.py -m gpu -x -q > gpurun_out/pytest_k1.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_k1.log
for v in default i2 i2m6; do
  if [ $v = default ]; then L=""; else L="$PWD/azurekinect3dreconstruction_b200/libb2recon_$v.so"; fi
  for m in 0 2; do
  B2R_K1_DBG=$m B2R_LIB=$L timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/k1var_${v}_$m.json 2> gpurun_out/k1var_${v}_$m.err
  done
done
B2R_LIB=$PWD/azurekinect3dreconstruction_b200/libb2recon_tr.so timeout 300 python tools/k1_trace.py > gpurun_out/k1_trace.log 2>&1
