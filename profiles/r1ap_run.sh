set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/r1ap_bench.json 2> gpurun_out/r1ap_bench.err; echo bench rc=$?
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1ap_bench_reference.json 2> gpurun_out/r1ap_ref.err; echo ref rc=$?
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r1ap_launches_bench_c3.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r1ap_ncu.log 2>&1; echo ncu rc=$?
for wl in c1 c2 c5; do timeout 300 python bench.py --workload $wl --no-cpu-baseline > gpurun_out/r1ap_bench_$wl.json 2> gpurun_out/r1ap_bench_$wl.err; echo $wl rc=$?; done
