#!/bin/bash
# FastBoxBlur: parity tests, C5 bench (packed kernel on / off), pipe utilisation of the packed kernel
cd /root/repo
O=gpurun_out; mkdir -p $O; T=${1:-r2f}
timeout 600 python -m pytest tests -m gpu -x -q -k "box" > $O/${T}_pytest.log 2>&1; tail -5 $O/${T}_pytest.log
for v in 1; do
  BLUR_B200_BOX_PACKED=$v timeout 300 python bench.py --workload box_c5 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > $O/${T}_box_$v.json 2> $O/${T}_box_$v.err; tail -1 $O/${T}_box_$v.json | cut -c1-400; tail -2 $O/${T}_box_$v.err
done
M="gpu__time_duration.sum,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum"
timeout 300 ncu --metrics $M --clock-control none -k regex:'box' -s 4 -c 1 --csv --log-file $O/${T}_pipes.csv \
   python bench.py --workload box_c5 --steps 2 --warmup 2 --no-cpu-baseline --no-e2e > $O/${T}_pipes.log 2>&1
grep -E "box" $O/${T}_pipes.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}' | tr -d '"'
