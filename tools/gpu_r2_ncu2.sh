mkdir -p gpurun_out
export GEOCB_SSSP_MAX_ROUNDS=100000
timeout 900 python tools/sweep_vm.py --blocks 1 --scales 4 --minact 16 --trace > gpurun_out/r2_sssp_round_trace.log 2>&1; tail -8 gpurun_out/r2_sssp_round_trace.log | cut -c1-400
CMD1="python tools/prof_edge_len.py 1310720"
CMD2="python tools/prof_knn.py 20000 16"
$CMD1 > gpurun_out/ncu_plain1.log 2>&1 && $CMD2 > gpurun_out/ncu_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemm3x_kernel -s 3 -c 1 -f -o gpurun_out/r2_gemm3x_prof $CMD1 > gpurun_out/ncu_gemm.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:knn_filter_tc_kernel -s 1 -c 1 -f -o gpurun_out/r2_knn_tc_prof $CMD2 > gpurun_out/ncu_knn.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_gemm.log; tail -3 gpurun_out/ncu_knn.log
