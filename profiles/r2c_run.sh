set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_umma.py -m gpu -q -x 2>&1 | tail -5 > gpurun_out/r2c_pytest.log; tail -3 gpurun_out/r2c_pytest.log
timeout 600 python bench.py --no-cpu-baseline --no-reference-gpu > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo bench rc=$?; tail -3 gpurun_out/r2c_bench.err
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r2c_launches_bench_c3.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-reference-gpu > gpurun_out/r2c_ncu.log 2>&1; echo ncu rc=$?
