set -x
mkdir -p gpurun_out
timeout 600 python profiles/graph_kernels.py c1 20 > gpurun_out/r3g_graph_kernels_c1.txt 2> gpurun_out/r3g_gk_c1.err; head -45 gpurun_out/r3g_graph_kernels_c1.txt
timeout 900 python bench.py --workload c5 --no-reference-gpu --no-cpu-baseline > gpurun_out/r3g_bench_c5.json 2> gpurun_out/r3g_bench_c5.err; tail -3 gpurun_out/r3g_bench_c5.err
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r3g_bench_*.json')):
    try:
        j=json.load(open(f)); print(f, round(j['value']), round(j['ms_per_step'],3), 'e2e', round(j['e2e']['value']), 'eval', round(j['eval_value']))
        for k,v in j['roofline']['all_kernels'].items(): print('  ',k, round(v['frac'],3), round(v['ms_per_step'],3))
    except Exception as e: print(f, 'ERR', e)
P
