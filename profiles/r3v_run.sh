set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -25 > gpurun_out/r3v_pytest.log; tail -4 gpurun_out/r3v_pytest.log
for w in c1 c3; do
  timeout 600 python bench.py --workload $w --steps 30 --no-reference-gpu --no-cpu-baseline > gpurun_out/r3v_bench_${w}.json 2> gpurun_out/r3v_bench_${w}.err; tail -2 gpurun_out/r3v_bench_${w}.err
done
timeout 600 python profiles/graph_kernels.py c1 20 > gpurun_out/r3v_graph_kernels_c1.txt 2> gpurun_out/r3v_gk_c1.err; grep -i "adam\|pos_grad\|reduce_partials\|kernel time" gpurun_out/r3v_graph_kernels_c1.txt
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r3v_bench_*.json')):
    try:
        j=json.load(open(f)); print(f, round(j['value']), round(j['ms_per_step'],3), 'e2e', round(j['e2e']['value']), 'eval', round(j['eval_value']))
        for k in ('filter_fwd','filter_bwd'): print('   ',k, round(j['roofline']['all_kernels'][k]['frac'],3), round(j['roofline']['all_kernels'][k]['ms_per_step'],4))
    except Exception as e: print(f, 'ERR', e)
P
