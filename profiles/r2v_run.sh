set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "duorec or sasrec" 2>&1 | tail -12 > gpurun_out/r2v_pytest.log; tail -8 gpurun_out/r2v_pytest.log
