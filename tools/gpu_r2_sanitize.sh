# usage: bash tools/gpu_r2_sanitize.sh memcheck|racecheck      (one tool per gpurun call)
TOOL=${1:-memcheck}
mkdir -p gpurun_out
export GEOCB_SSSP_MAX_ROUNDS=100000
timeout 300 python tools/sanitize_target.py > gpurun_out/r2_sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/r2_sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 30 python tools/sanitize_target.py > gpurun_out/r2_sanitizer_$TOOL.txt 2>&1
echo "sanitizer rc=$?"; tail -15 gpurun_out/r2_sanitizer_$TOOL.txt
