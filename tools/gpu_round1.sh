set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err; echo "bench rc=$?" >> gpurun_out/bench_a.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:tsdf -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1 ; \
ncu --set full --clock-control none --import-source on -k regex:tsdf_fuse -s 100 -c 3 -o gpurun_out/prof_fuse python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1 ; \
ncu --set full --clock-control none --import-source on -k regex:tsdf_pack -s 100 -c 2 -o gpurun_out/prof_pack python bench.py --steps 1 --warmup 1 --frames 24 --no-cpu-baseline > gpurun_out/ncu3.log 2>&1
nvidia-smi > gpurun_out/nvsmi.txt; lscpu | head -20 > gpurun_out/lscpu.txt; nproc >> gpurun_out/lscpu.txt
