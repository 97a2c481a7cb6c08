set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -12 > gpurun_out/r2e_pytest.log; tail -6 gpurun_out/r2e_pytest.log
