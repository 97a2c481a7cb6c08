set -x
mkdir -p gpurun_out
BSA_E_TRUNC=1 timeout 900 python -m pytest tests/test_gpu_umma.py -m gpu -q -x -k "cross_entropy or ce_backward" 2>&1 | tail -15 > gpurun_out/r3d_pytest_trunc.log; tail -12 gpurun_out/r3d_pytest_trunc.log
BSA_E_TRUNC=1 timeout 600 python bench.py --no-reference-gpu --no-cpu-baseline > gpurun_out/r3d_bench_c3_trunc.json 2> gpurun_out/r3d_bench_c3_trunc.err; tail -3 gpurun_out/r3d_bench_c3_trunc.err
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r3d_bench_*.json')):
    try:
        j=json.load(open(f)); print(f, round(j['value']), round(j['ms_per_step'],3), 'e2e', round(j['e2e']['value']), 'eval', round(j['eval_value']))
        for k,v in j['roofline']['all_kernels'].items(): print('  ',k, round(v['frac'],3), round(v['ms_per_step'],3))
    except Exception as e: print(f, 'ERR', e)
P
