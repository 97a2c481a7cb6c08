set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > gpurun_out/r4g_pytest.log; tail -3 gpurun_out/r4g_pytest.log
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -2
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/r4g_bench.json 2> gpurun_out/r4g_bench.err; echo bench rc=$?; tail -2 gpurun_out/r4g_bench.err
python - <<'P'
import json
j=json.load(open('gpurun_out/r4g_bench.json')); print(round(j['value']), round(j['ms_per_step'],3), 'e2e', round(j['e2e']['value']), 'eval', round(j['eval_value']), j['roofline']['kernel'], round(j['roofline']['frac'],3), j['roofline']['traffic'], j['clocks'])
P
