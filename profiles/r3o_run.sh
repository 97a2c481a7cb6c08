set -x
mkdir -p gpurun_out
python profiles/attn_probe.py > gpurun_out/r3o_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_ --launch-skip 3 --launch-count 3 -o gpurun_out/r3o_attn python profiles/attn_probe.py > gpurun_out/r3o_ncu.log 2>&1; echo ncu rc=$?
ls -la gpurun_out/r3o_attn.ncu-rep
