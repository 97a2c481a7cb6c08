set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -25 > gpurun_out/r3p_pytest.log; tail -8 gpurun_out/r3p_pytest.log
timeout 600 python bench.py --no-reference-gpu --no-cpu-baseline > gpurun_out/r3p_bench_c3.json 2> gpurun_out/r3p_bench_c3.err; tail -3 gpurun_out/r3p_bench_c3.err
for w in c1 c2; do
  timeout 600 python bench.py --workload $w --steps 30 --no-reference-gpu --no-cpu-baseline > gpurun_out/r3p_bench_${w}_graph.json 2> gpurun_out/r3p_bench_${w}_graph.err; tail -3 gpurun_out/r3p_bench_${w}_graph.err
done
timeout 600 python profiles/step_breakdown.py c3 3 > gpurun_out/r3p_step_breakdown.txt 2> gpurun_out/r3p_sb.err; echo sb rc=$?
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r3p_bench_*.json')):
    try:
        j=json.load(open(f)); print(f, round(j['value']), round(j['ms_per_step'],3), 'e2e', round(j['e2e']['value']), 'eval', round(j['eval_value']), 'eval_e2e', round(j['eval_e2e']['value']), j['gpu_launches'])
    except Exception as e: print(f, 'ERR', e)
P
head -40 gpurun_out/r3p_step_breakdown.txt
