#!/bin/bash
# ncu --set full of the FFT kernels on the headline config.  usage: bash tools/gpu_ncu.sh TAG [kernel-regex]
TAG=${1:-r2}
RE=${2:-'ols2?_(row|col)_kernel|interleave'}
O=gpurun_out; mkdir -p $O
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $O/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$RE" -s 6 -c 3 -f -o $O/${TAG}_full \
   python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $O/${TAG}_ncu_full.log 2>&1
tail -2 $O/${TAG}_plain.log; tail -3 $O/${TAG}_ncu_full.log
