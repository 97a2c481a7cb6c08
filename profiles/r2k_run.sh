set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tests/dp_selfcheck.py > gpurun_out/r2k_dp_selfcheck.log 2>&1; echo selfcheck rc=$?; grep -n "vp vs bucket\|selfcheck ok\|AssertionError" gpurun_out/r2k_dp_selfcheck.log | head -40
