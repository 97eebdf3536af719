#!/bin/bash
# The one GPU-box runner (under gpurun):   bash tools/gpu_run.sh <step> [<step> ...]
# Every step writes its log under gpurun_out/ and never stops the ones after it.  Steps:
#   info            nvidia-smi, host cores and memory
#   tests[:EXPR]    pytest -m gpu (optionally -k EXPR)
#   smoke           __graft_entry__.smoke()
#   bench[:ARGS]    python bench.py ARGS                      -> bench.json / bench.err
#   ref[:ARGS]      python bench.py --impl reference ARGS     -> bench_ref.json
#   benchN:N[:ARGS] torchrun --nproc-per-node N bench.py --gpus N ARGS -> bench_nN.json
#   ops[:ARGS]      python tools/bench_ops.py ARGS            -> bench_ops.json
#   probe[:ARGS]    python tools/shard_probe.py ARGS          -> shard_probe.json
#   launches[:ARGS] ncu launch list of bench.py ARGS (after a plain run)   -> launches.csv
#   ncu:KERNEL[:ARGS]  ncu --set full of kernel regex KERNEL in bench.py ARGS -> prof_KERNEL.ncu-rep
#   py:SCRIPT[:ARGS]   python SCRIPT ARGS -> <basename>.json
set -u
mkdir -p gpurun_out
OUT=gpurun_out
SHORT="--steps 2 --warmup 3 --no-cpu-baseline --no-secondary"
for step in "$@"; do
  name="${step%%:*}"; rest=""; [[ "$step" == *:* ]] && rest="${step#*:}"
  echo "== $step" >&2
  case "$name" in
    info) { nvidia-smi; nproc; free -g; } > $OUT/info.log 2>&1 ;;
    tests) timeout 1500 python -m pytest tests -m gpu -x -q ${rest:+-k "$rest"} > $OUT/pytest_gpu.log 2>&1; echo "rc=$?" >> $OUT/pytest_gpu.log; tail -5 $OUT/pytest_gpu.log >&2 ;;
    smoke) timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "rc=$?" >> $OUT/smoke.log; tail -2 $OUT/smoke.log >&2 ;;
    bench) timeout 900 python bench.py $rest > $OUT/bench.json 2> $OUT/bench.err; echo "rc=$?" >> $OUT/bench.err; tail -3 $OUT/bench.err >&2 ;;
    ref) timeout 1500 python bench.py --impl reference $rest > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "rc=$?" >> $OUT/bench_ref.err ;;
    benchN) n="${rest%%:*}"; a=""; [[ "$rest" == *:* ]] && a="${rest#*:}"
            timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n $a \
              > $OUT/bench_n$n.json 2> $OUT/bench_n$n.err; echo "rc=$?" >> $OUT/bench_n$n.err; tail -3 $OUT/bench_n$n.err >&2 ;;
    ops) timeout 900 python tools/bench_ops.py $rest > $OUT/bench_ops.json 2> $OUT/bench_ops.err; echo "rc=$?" >> $OUT/bench_ops.err ;;
    probe) timeout 900 python tools/shard_probe.py $rest > $OUT/shard_probe.json 2> $OUT/shard_probe.err; echo "rc=$?" >> $OUT/shard_probe.err ;;
    launches) a="${rest:-$SHORT}"
            timeout 600 python bench.py $a > $OUT/plain.log 2>&1 && \
            timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $OUT/launches.csv python bench.py $a > $OUT/ncu_launches.log 2>&1
            echo "rc=$?" >> $OUT/ncu_launches.log ;;
    ncu) k="${rest%%:*}"; a="$SHORT"; [[ "$rest" == *:* ]] && a="${rest#*:}"
            timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$k" -s 4 -c 2 -f -o $OUT/prof_$k python bench.py $a > $OUT/ncu_$k.log 2>&1
            echo "rc=$?" >> $OUT/ncu_$k.log ;;
    py) sc="${rest%%:*}"; a=""; [[ "$rest" == *:* ]] && a="${rest#*:}"; b=$(basename "$sc" .py)
            timeout 1200 python $sc $a > $OUT/$b.json 2> $OUT/$b.err; echo "rc=$?" >> $OUT/$b.err; tail -3 $OUT/$b.err >&2 ;;
    *) echo "unknown step $step" >&2 ;;
  esac
done
exit 0
