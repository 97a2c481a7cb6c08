set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r3l_bench.json 2> gpurun_out/r3l_bench.err; echo bench rc=$?; tail -3 gpurun_out/r3l_bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r3l_bench_reference.json 2> gpurun_out/r3l_ref.err; echo ref rc=$?
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --steps 2 --warmup 3 --graph off --no-cpu-baseline --no-reference-gpu > gpurun_out/r3l_plain.json 2> gpurun_out/r3l_plain.err && timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r3l_launches_bench_c3.csv python bench.py --steps 2 --warmup 3 --graph off --no-cpu-baseline --no-reference-gpu > gpurun_out/r3l_ncu.log 2>&1; echo ncu rc=$?
python - <<'P'
import json
j=json.load(open('gpurun_out/r3l_bench.json')); print(round(j['value']), round(j['ms_per_step'],3), 'e2e', round(j['e2e']['value']), 'eval', round(j['eval_value']), j.get('reference_gpu'), j.get('cpu_baseline'))
P
