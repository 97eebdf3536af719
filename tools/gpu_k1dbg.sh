set -x
for m in 1 3 4; do
B2R_K1_DBG=$m timeout 600 python bench.py --steps 3 --no-cpu-baseline > gpurun_out/bench_k1dbg$m.json 2> gpurun_out/bench_k1dbg$m.err
done
