set -x
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_golden.py tests/test_gpu_fullsize.py -m gpu -x -q > gpurun_out/pytest_k1.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_k1.log
for m in 0 1 2; do
B2R_K1_DBG=$m timeout 600 python bench.py --steps 3 --no-cpu-baseline > gpurun_out/bench_k1dbg$m.json 2> gpurun_out/bench_k1dbg$m.err
done
B2R_K1_TMA=0 timeout 600 python bench.py --steps 3 --no-cpu-baseline > gpurun_out/bench_k1dbg9.json 2> gpurun_out/bench_k1dbg9.err
