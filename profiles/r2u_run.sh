set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/r2u_pytest.log; tail -3 gpurun_out/r2u_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2u_bench.json 2> gpurun_out/r2u_bench.err; echo bench rc=$?; tail -3 gpurun_out/r2u_bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r2u_bench_reference.json 2> gpurun_out/r2u_ref.err; echo ref rc=$?
for wl in c1 c2; do timeout 300 python bench.py --workload $wl --no-cpu-baseline --no-reference-gpu > gpurun_out/r2u_bench_$wl.json 2> gpurun_out/r2u_bench_$wl.err; echo $wl rc=$?; done
timeout 600 python profiles/step_breakdown.py c3 3 > gpurun_out/r2u_step_breakdown.txt 2> gpurun_out/r2u_sb.err; echo sb rc=$?
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-reference-gpu > gpurun_out/r2u_plain.json 2> gpurun_out/r2u_plain.err && timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r2u_launches_bench_c3.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-reference-gpu > gpurun_out/r2u_ncu.log 2>&1; echo ncu rc=$?
