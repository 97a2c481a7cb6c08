set -x
mkdir -p gpurun_out
for w in c2 c5; do
timeout 600 python bench.py --workload $w --steps 30 --no-reference-gpu --no-cpu-baseline > gpurun_out/r4d_bench_${w}_side.json 2> gpurun_out/r4d_${w}_side.err
BSA_SIDE_STREAM=0 timeout 600 python bench.py --workload $w --steps 30 --no-reference-gpu --no-cpu-baseline > gpurun_out/r4d_bench_${w}_base.json 2> gpurun_out/r4d_${w}_base.err
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r4d_bench_*.json')):
    try:
        j=json.load(open(f)); print(f, round(j['value']), round(j['ms_per_step'],3), 'e2e', round(j['e2e']['value']), 'eval', round(j['eval_value']))
    except Exception as e: print(f, 'ERR', e)
P
