set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/r1aj_bench.json 2> gpurun_out/r1aj_bench.err; echo bench rc=$?
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1aj_bench_reference.json 2> gpurun_out/r1aj_ref.err; echo ref rc=$?
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r1aj_launches_bench_c3.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r1aj_ncu.log 2>&1; echo ncu rc=$?
timeout 300 python profiles/step_breakdown.py > gpurun_out/r1aj_step_breakdown.txt 2> gpurun_out/r1aj_sb.err; echo sb rc=$?
BSA_EXTRA_NVCC_FLAGS=-DBSA_ATTN_STAMPS timeout 600 python profiles/attn_phases.py > gpurun_out/r1aj_attn_phases.txt 2> gpurun_out/r1aj_ph.err; echo phases rc=$?
