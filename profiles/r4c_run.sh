set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > gpurun_out/r4c_pytest.log; tail -3 gpurun_out/r4c_pytest.log
for i in 1 2; do
for w in c1 c3; do
timeout 600 python bench.py --workload $w --steps 30 --no-reference-gpu --no-cpu-baseline > gpurun_out/r4c_bench_${w}_side$i.json 2> gpurun_out/r4c_${w}_side$i.err
BSA_SIDE_STREAM=0 timeout 600 python bench.py --workload $w --steps 30 --no-reference-gpu --no-cpu-baseline > gpurun_out/r4c_bench_${w}_base$i.json 2> gpurun_out/r4c_${w}_base$i.err
done; done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r4c_bench_*.json')):
    try:
        j=json.load(open(f)); print(f, round(j['value']), round(j['ms_per_step'],3), 'e2e', round(j['e2e']['value']), 'eval', round(j['eval_value']))
    except Exception as e: print(f, 'ERR', e)
P
