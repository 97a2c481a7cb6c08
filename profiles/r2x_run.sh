set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -40 > gpurun_out/r2x_pytest.log; tail -30 gpurun_out/r2x_pytest.log
for w in c1 c2; do
  timeout 600 python bench.py --workload $w --graph on --steps 30 --no-reference-gpu > gpurun_out/r2x_bench_${w}_graph.json 2> gpurun_out/r2x_bench_${w}_graph.err; tail -3 gpurun_out/r2x_bench_${w}_graph.err
  timeout 600 python bench.py --workload $w --graph off --steps 30 --no-reference-gpu --no-cpu-baseline > gpurun_out/r2x_bench_${w}_eager.json 2> gpurun_out/r2x_bench_${w}_eager.err
done
timeout 900 python bench.py --no-reference-gpu --no-cpu-baseline > gpurun_out/r2x_bench_c3.json 2> gpurun_out/r2x_bench_c3.err; tail -3 gpurun_out/r2x_bench_c3.err
timeout 900 python bench.py --graph on --no-reference-gpu --no-cpu-baseline > gpurun_out/r2x_bench_c3_graph.json 2> gpurun_out/r2x_bench_c3_graph.err; tail -3 gpurun_out/r2x_bench_c3_graph.err
timeout 900 python bench.py --workload c5 --graph on --no-reference-gpu --no-cpu-baseline > gpurun_out/r2x_bench_c5_graph.json 2> gpurun_out/r2x_bench_c5_graph.err; tail -3 gpurun_out/r2x_bench_c5_graph.err
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2x_bench_*.json')):
    try:
        j=json.load(open(f)); print(f, round(j['value']), round(j['ms_per_step'],3), 'e2e', round(j['e2e']['value']), 'eval', round(j['eval_value']), 'eval_e2e', round(j['eval_e2e']['value']), j['gpu_launches'])
    except Exception as e: print(f, 'ERR', e)
P
