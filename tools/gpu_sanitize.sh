#!/bin/bash
# one compute-sanitizer tool per gpurun call (B200_PROFILING.md): tools/gpu_sanitize.sh racecheck|synccheck|memcheck
cd "$(dirname "$0")/.." || exit 1
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
tool=$1
O=gpurun_out
mkdir -p $O
timeout 600 python tools/sanitize_target.py > $O/r2_sanitize_plain_$tool.log 2>&1 || { echo "plain run failed"; tail -5 $O/r2_sanitize_plain_$tool.log; exit 1; }
timeout 2400 compute-sanitizer --tool $tool --print-limit 40 --log-file $O/r2_sanitize_$tool.log python tools/sanitize_target.py > $O/r2_sanitize_${tool}_stdout.log 2>&1
echo "rc $?" >> $O/r2_sanitize_${tool}_stdout.log
tail -15 $O/r2_sanitize_$tool.log
