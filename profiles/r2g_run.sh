set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_named_configs.py -m gpu -q -x 2>&1 | tail -15 > gpurun_out/r2g_pytest.log; tail -6 gpurun_out/r2g_pytest.log
timeout 600 python bench.py --no-cpu-baseline --no-reference-gpu > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err; echo bench rc=$?; tail -3 gpurun_out/r2g_bench.err
BSA_ATTN_BWD_UMMA=1 timeout 600 python bench.py --no-cpu-baseline --no-reference-gpu > gpurun_out/r2g_bench_kvsync.json 2> gpurun_out/r2g_bench2.err; echo bench2 rc=$?
