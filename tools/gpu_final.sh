set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python tools/config5.py > gpurun_out/config5.json 2> gpurun_out/config5.err; echo "rc=$?" >> gpurun_out/config5.err
