# usage: bash profiles/n2exp.sh N   — NCCL protocol / channel experiments for the data-parallel train step
N=${1:-2}
i=0
for cfg in "" "NCCL_PROTO=Simple" "NCCL_PROTO=Simple NCCL_MAX_CTAS=16" "NCCL_PROTO=LL128" "NCCL_ALGO=NVLS" ; do
  i=$((i+1))
  env $cfg python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29540+i)) bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_u_n${N}_$i.json 2> gpurun_out/bench_u_n${N}_$i.err
  python -c "
import json,sys
try:
    d=json.load(open('gpurun_out/bench_u_n${N}_$i.json')); print('N=$N [$cfg]', round(d['value']), round(d['ms_per_step'],3), round(d['eval']['value']))
except Exception as e: print('N=$N [$cfg] failed', e)"
done
