set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_named_configs.py tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -4 > gpurun_out/r2t_pytest.log; tail -2 gpurun_out/r2t_pytest.log
timeout 300 python bench.py --workload c5 --no-cpu-baseline --no-reference-gpu > gpurun_out/r2t_bench_c5.json 2> gpurun_out/r2t_c5.err; echo c5 rc=$?; tail -2 gpurun_out/r2t_c5.err
BSA_ATTN_BWD_UMMA=1 timeout 300 python bench.py --workload c5 --no-cpu-baseline --no-reference-gpu > gpurun_out/r2t_bench_c5_sync.json 2> gpurun_out/r2t_c5b.err; echo c5b rc=$?
