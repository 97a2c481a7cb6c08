set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > gpurun_out/r4h_pytest.log; tail -3 gpurun_out/r4h_pytest.log
for i in 1 2; do
timeout 600 python bench.py --no-reference-gpu --no-cpu-baseline > gpurun_out/r4h_bench_c3_unstaged$i.json 2> gpurun_out/r4h_u$i.err
BSA_FILTER_STAGE_KB=220 timeout 600 python bench.py --no-reference-gpu --no-cpu-baseline > gpurun_out/r4h_bench_c3_staged$i.json 2> gpurun_out/r4h_s$i.err
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r4h_bench_*.json')):
    try:
        j=json.load(open(f)); a=j['roofline']['all_kernels']; print(f, round(j['value']), round(j['ms_per_step'],3), 'eval', round(j['eval_value']), 'filter fwd/bwd ms', round(a['filter_fwd']['ms_per_step'],3), round(a['filter_bwd']['ms_per_step'],3))
    except Exception as e: print(f, 'ERR', e)
P
