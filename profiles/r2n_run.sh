set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_named_configs.py tests/test_gpu_data.py -m gpu -q -x 2>&1 | tail -6 > gpurun_out/r2n_pytest.log; tail -3 gpurun_out/r2n_pytest.log
timeout 600 python bench.py --no-cpu-baseline --no-reference-gpu > gpurun_out/r2n_bench.json 2> gpurun_out/r2n_bench.err; echo bench rc=$?; tail -3 gpurun_out/r2n_bench.err
timeout 300 python bench.py --workload c1 --no-cpu-baseline --no-reference-gpu > gpurun_out/r2n_bench_c1.json 2> gpurun_out/r2n_bench_c1.err; echo c1 rc=$?
timeout 300 python bench.py --workload c2 --no-cpu-baseline --no-reference-gpu > gpurun_out/r2n_bench_c2.json 2> gpurun_out/r2n_bench_c2.err; echo c2 rc=$?
