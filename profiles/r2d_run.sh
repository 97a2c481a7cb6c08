set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $TR bench.py --gpus 2 > gpurun_out/r2d_bench_n2.json 2> gpurun_out/r2d_bench_n2.err; echo n2 rc=$?; tail -5 gpurun_out/r2d_bench_n2.err
timeout 900 $TR bench.py --gpus 2 --workload c4 --users 1000000 > gpurun_out/r2d_bench_c4_n2.json 2> gpurun_out/r2d_c4_n2.err; echo c4 rc=$?; tail -5 gpurun_out/r2d_c4_n2.err
timeout 600 $TR tests/dp_selfcheck.py > gpurun_out/r2d_dp_selfcheck.log 2>&1; echo selfcheck rc=$?; tail -3 gpurun_out/r2d_dp_selfcheck.log
