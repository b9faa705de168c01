export GEOCB_SSSP_MAX_ROUNDS=100000
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sssp_vm.py -x -q > gpurun_out/r2_vm_tests.log 2>&1; echo "vm tests rc=$?" >> gpurun_out/r2_vm_tests.log
tail -15 gpurun_out/r2_vm_tests.log
timeout 900 python tools/sweep_vm.py --blocks 1,8,15,24 --scales 4 --minact 16 --trace > gpurun_out/r2_sweep_vm_1.log 2>&1; echo "sweep rc=$?" >> gpurun_out/r2_sweep_vm_1.log
tail -30 gpurun_out/r2_sweep_vm_1.log
