set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_umma.py tests/test_gpu_index_exact.py -m gpu -q -x 2>&1 | tail -15 > gpurun_out/r2b_pytest.log; tail -5 gpurun_out/r2b_pytest.log
timeout 600 python -m pytest tests/test_gpu_named_configs.py tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -15 > gpurun_out/r2b_pytest2.log; tail -5 gpurun_out/r2b_pytest2.log
timeout 600 python bench.py --no-cpu-baseline --no-reference-gpu > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo bench rc=$?; tail -3 gpurun_out/r2b_bench.err
timeout 600 python bench.py --workload c4 --users 300000 --no-cpu-baseline > gpurun_out/r2b_bench_c4_n1.json 2> gpurun_out/r2b_c4.err; echo c4 rc=$?; tail -3 gpurun_out/r2b_c4.err
