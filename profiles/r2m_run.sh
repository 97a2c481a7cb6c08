set -x
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-reference-gpu > gpurun_out/r2m_plain.json 2> gpurun_out/r2m_plain.err && ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"EpiLse" -s 4 -c 1 -o gpurun_out/r2m_celse python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-reference-gpu > gpurun_out/r2m_ncu.log 2>&1; echo ncu rc=$?
tail -3 gpurun_out/r2m_ncu.log
