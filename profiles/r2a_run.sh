set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/r2a_pytest.log; tail -3 gpurun_out/r2a_pytest.log
timeout 900 python -m pytest tests/test_gpu_reference_trainer_dropin.py -m gpu -v 2>&1 | tail -15 > gpurun_out/r2a_dropin_reference_trainer.log; tail -4 gpurun_out/r2a_dropin_reference_trainer.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo bench rc=$?; tail -3 gpurun_out/r2a_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2a_bench_reference.json 2> gpurun_out/r2a_ref.err; echo ref rc=$?
timeout 600 python bench.py --workload c4 --users 300000 --no-cpu-baseline > gpurun_out/r2a_bench_c4_n1.json 2> gpurun_out/r2a_c4.err; echo c4 rc=$?; tail -3 gpurun_out/r2a_c4.err
for wl in c1 c2 c5; do timeout 300 python bench.py --workload $wl --no-cpu-baseline --no-reference-gpu > gpurun_out/r2a_bench_$wl.json 2> gpurun_out/r2a_bench_$wl.err; echo $wl rc=$?; done
