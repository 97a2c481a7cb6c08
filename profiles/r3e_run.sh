set -x
mkdir -p gpurun_out
for i in 1 2; do
BSA_E_TRUNC=1 timeout 600 python bench.py --no-reference-gpu --no-cpu-baseline > gpurun_out/r3e_bench_c3_trunc$i.json 2> gpurun_out/r3e_bench_c3_trunc$i.err
BSA_E_TRUNC=0 timeout 600 python bench.py --no-reference-gpu --no-cpu-baseline > gpurun_out/r3e_bench_c3_base$i.json 2> gpurun_out/r3e_bench_c3_base$i.err
done
BSA_E_TRUNC=1 timeout 600 python profiles/step_breakdown.py c3 3 2>/dev/null | grep -i "split\|EpiLse\|ce_bwd\|sum of kernel" | head
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r3e_bench_*.json')):
    try:
        j=json.load(open(f)); print(f, round(j['value']), round(j['ms_per_step'],3), 'e2e', round(j['e2e']['value']), 'eval', round(j['eval_value']))
    except Exception as e: print(f, 'ERR', e)
P
