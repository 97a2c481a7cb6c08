set -x
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tests/dp_selfcheck.py > gpurun_out/r4e_dp_selfcheck_n$N.log 2>&1; echo selfcheck rc=$?; grep -n "vp vs bucket\|selfcheck ok\|AssertionError\|Error" gpurun_out/r4e_dp_selfcheck_n$N.log | head -8 | cut -c1-400
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r4e_bench_n$N.json 2> gpurun_out/r4e_bench_n$N.err; echo bench rc=$?; tail -3 gpurun_out/r4e_bench_n$N.err
