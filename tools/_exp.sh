set -x
timeout 600 python -m pytest tests/test_reftoys.py tests/test_host_api.py -m gpu -q 2>&1 | tail -12 > gpurun_out/exp8_toys.log
