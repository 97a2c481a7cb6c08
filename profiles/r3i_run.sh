set -x
mkdir -p gpurun_out
python profiles/ncu_train_steps.py 3 > gpurun_out/r3i_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:umma_gemm_kernel --launch-skip 28 --launch-count 5 -o gpurun_out/r3i_ffn_gemms python profiles/ncu_train_steps.py 3 > gpurun_out/r3i_ncu.log 2>&1; echo ncu rc=$?
ls -la gpurun_out/r3i_ffn_gemms.ncu-rep
