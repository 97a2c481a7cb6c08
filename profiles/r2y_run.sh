set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -15 > gpurun_out/r2y_pytest_no_alloc_alone.log; cat gpurun_out/r2y_pytest_no_alloc_alone.log
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -25 > gpurun_out/r2y_pytest.log; tail -8 gpurun_out/r2y_pytest.log
timeout 600 python profiles/graph_kernels.py c1 20 > gpurun_out/r2y_graph_kernels_c1.txt 2> gpurun_out/r2y_gk_c1.err; tail -3 gpurun_out/r2y_gk_c1.err; head -12 gpurun_out/r2y_graph_kernels_c1.txt
timeout 600 python profiles/graph_kernels.py c2 20 > gpurun_out/r2y_graph_kernels_c2.txt 2> gpurun_out/r2y_gk_c2.err; head -12 gpurun_out/r2y_graph_kernels_c2.txt
