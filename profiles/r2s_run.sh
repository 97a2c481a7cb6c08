set -x
mkdir -p gpurun_out
timeout 600 python bench.py --no-cpu-baseline --no-reference-gpu > gpurun_out/r2s_bench_stage.json 2> gpurun_out/r2s_a.err; echo rc=$?
BSA_FILTER_STAGE_KB=0 timeout 600 python bench.py --no-cpu-baseline --no-reference-gpu > gpurun_out/r2s_bench_nostage.json 2> gpurun_out/r2s_b.err; echo rc=$?
BSA_FILTER_STAGE_KB=0 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -3
