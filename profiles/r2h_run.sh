set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -4 > gpurun_out/r2h_pytest.log; tail -2 gpurun_out/r2h_pytest.log
timeout 600 python bench.py --no-cpu-baseline --no-reference-gpu > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err; echo bench rc=$?; tail -3 gpurun_out/r2h_bench.err
timeout 600 python profiles/step_breakdown.py c3 3 > gpurun_out/r2h_step_breakdown.txt 2> gpurun_out/r2h_sb.err; echo sb rc=$?
