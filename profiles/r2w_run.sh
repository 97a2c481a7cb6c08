set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -25 > gpurun_out/r2w_pytest.log; tail -12 gpurun_out/r2w_pytest.log
