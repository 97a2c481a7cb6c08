set -x
mkdir -p gpurun_out
python profiles/ce_probe.py > gpurun_out/r2f_ce_probe.log 2>&1 && MODES=2 ITERS=1 TIME=0 ncu --set full --clock-control none --import-source on -k regex:"ce_bwd_logits|EpiLse" -c 3 -o gpurun_out/r2f_ce python profiles/ce_probe.py > gpurun_out/r2f_ncu.log 2>&1; echo ncu rc=$?
cat gpurun_out/r2f_ce_probe.log
