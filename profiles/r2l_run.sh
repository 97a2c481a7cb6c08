set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tests/dp_selfcheck.py > gpurun_out/r2l_dp_selfcheck_n$1.log 2>&1; echo selfcheck rc=$?; grep -n "vp vs bucket\|selfcheck ok\|AssertionError" gpurun_out/r2l_dp_selfcheck_n$1.log | head
timeout 900 $TR bench.py --gpus $1 --dp vp --steps 20 --warmup 5 > gpurun_out/r2l_bench_n$1_vp.json 2> gpurun_out/r2l_bench_n$1_vp.err; echo vp rc=$?; tail -3 gpurun_out/r2l_bench_n$1_vp.err
