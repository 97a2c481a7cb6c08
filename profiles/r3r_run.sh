set -x
mkdir -p gpurun_out
timeout 600 python profiles/graph_kernels.py c3 10 > gpurun_out/r3r_graph_kernels_c3.txt 2> gpurun_out/r3r_gk_c3.err; head -30 gpurun_out/r3r_graph_kernels_c3.txt; tail -3 gpurun_out/r3r_gk_c3.err
