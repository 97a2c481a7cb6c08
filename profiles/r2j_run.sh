set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_named_configs.py -m gpu -q -x 2>&1 | tail -6 > gpurun_out/r2j_pytest.log; tail -3 gpurun_out/r2j_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tests/dp_selfcheck.py > gpurun_out/r2j_dp_selfcheck.log 2>&1; echo selfcheck rc=$?; tail -5 gpurun_out/r2j_dp_selfcheck.log
timeout 900 $TR bench.py --gpus 2 --dp vp > gpurun_out/r2j_bench_n2_vp.json 2> gpurun_out/r2j_bench_n2_vp.err; echo vp rc=$?; tail -5 gpurun_out/r2j_bench_n2_vp.err
timeout 900 $TR bench.py --gpus 2 > gpurun_out/r2j_bench_n2.json 2> gpurun_out/r2j_bench_n2.err; echo n2 rc=$?
