export GEOCB_SSSP_MAX_ROUNDS=100000
mkdir -p gpurun_out
CMD="python tools/sweep_vm.py --blocks ${VM_B:-1} --scales 4 --minact 16 --skip-slot"
$CMD > gpurun_out/ncu_vm_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:sssp_vm_kernel -s 5 -c 1 -f -o gpurun_out/r2_vm_prof $CMD > gpurun_out/ncu_vm.log 2>&1
echo "rc=$?"; tail -5 gpurun_out/ncu_vm_plain.log; tail -5 gpurun_out/ncu_vm.log
