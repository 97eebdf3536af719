set -x
timeout 900 python -m pytest tests/test_gpu_fullsize.py tests/test_golden.py -m gpu -x -q > gpurun_out/pytest_new.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_new.log
