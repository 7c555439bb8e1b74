#!/bin/bash
# column walk reordered for 128-byte line sharing: timings, DRAM traffic, full GPU test-suite, bench
cd /root/repo
O=gpurun_out; mkdir -p $O; T=r2e
timeout 300 python tools/diag/pass_time.py --variants "OLS2=1" 2>&1 | tail -1
timeout 300 python tools/diag/pass_time.py --variants "OLS2=1" --sigma 5 2>&1 | tail -1
timeout 900 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1; tail -5 $O/${T}_pytest.log
for wl in gauss_c3 gauss_c2 gauss_c1 box_c5; do
  timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 > $O/${T}_bench_$wl.json 2> $O/${T}_bench_$wl.err; tail -1 $O/${T}_bench_$wl.json; tail -2 $O/${T}_bench_$wl.err
done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err; tail -1 $O/${T}_bench_ref.json; tail -2 $O/${T}_bench_ref.err
M="dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct"
timeout 300 ncu --metrics $M --clock-control none -k regex:'ols2_' -s 4 -c 2 --csv --log-file $O/${T}_dram.csv \
      python tools/diag/pass_time.py --reps 1 --variants "OLS2=1" > $O/${T}_dram.log 2>&1
grep -E "ols2_" $O/${T}_dram.csv | awk -F'","' '{print substr($5,1,30), $(NF-2), $(NF-1), $NF}' | tr -d '"'
