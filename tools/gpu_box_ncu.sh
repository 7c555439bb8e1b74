#!/bin/bash
cd /root/repo
O=gpurun_out; mkdir -p $O; T=${1:-r2g}
timeout 300 python bench.py --workload box_c5 --steps 2 --warmup 2 --no-cpu-baseline --no-e2e > $O/${T}_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"box" -s 4 -c 2 -f -o $O/${T}_box_full \
   python bench.py --workload box_c5 --steps 2 --warmup 2 --no-cpu-baseline --no-e2e > $O/${T}_ncu.log 2>&1
tail -1 $O/${T}_plain.log | cut -c1-300; tail -3 $O/${T}_ncu.log
