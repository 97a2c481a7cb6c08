set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/r2p_pytest.log; tail -3 gpurun_out/r2p_pytest.log
timeout 300 python profiles/ce_probe.py > gpurun_out/r2p_ce_probe.log 2>&1; cat gpurun_out/r2p_ce_probe.log
timeout 600 python bench.py --no-cpu-baseline --no-reference-gpu > gpurun_out/r2p_bench.json 2> gpurun_out/r2p_bench.err; echo bench rc=$?; tail -3 gpurun_out/r2p_bench.err
