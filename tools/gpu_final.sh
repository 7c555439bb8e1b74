#!/bin/bash
# Final GPU-box visit of a round: gpu tests, smoke, bench lines of every workload (+ pffft mode) and the reference arm,
# the other configs (C4 ...), the 45-size sweep (flags 2-5) for the plot, ncu launch lists and full captures.
# usage (from the repo root, under gpurun): bash tools/gpu_final.sh TAG
TAG=${1:-r02z}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1; nproc >> $O/${TAG}_smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> $O/${TAG}_smoke.log
for wl in gauss_c3 gauss_c2 gauss_c1 box_c5; do
  timeout 600 python bench.py --workload $wl --steps 20 --warmup 3 > $O/${TAG}_bench_$wl.json 2> $O/${TAG}_bench_$wl.err; echo "bench rc=$?" >> $O/${TAG}_bench_$wl.err
done
timeout 300 python bench.py --nyquist pffft --steps 20 --warmup 3 --no-cpu-baseline --no-e2e > $O/${TAG}_bench_gauss_c3_pffft.json 2> $O/${TAG}_bench_gauss_c3_pffft.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err
timeout 300 python tools/bench_configs.py > $O/${TAG}_configs.jsonl 2> $O/${TAG}_configs.err
for f in 3 5 2 4; do
  timeout 300 python -m blur_algorithms_b200.cli $f 0 synthetic --sweep > $O/${TAG}_sweep_flag$f.jsonl 2> $O/${TAG}_sweep_flag$f.err
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/${TAG}_launches.csv \
   python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-verify > $O/${TAG}_ncu_launch.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/${TAG}_launches_box.csv \
   python bench.py --workload box_c5 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-verify > $O/${TAG}_ncu_launch_box.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'ols2?_(row|col)_kernel|interleave' -s 6 -c 3 -f -o $O/${TAG}_full \
   python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-verify > $O/${TAG}_ncu_full.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'box' -s 4 -c 1 -f -o $O/${TAG}_box_full \
   python bench.py --workload box_c5 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-verify > $O/${TAG}_ncu_box_full.log 2>&1
tail -3 $O/${TAG}_pytest.log $O/${TAG}_smoke.log 2>/dev/null
for wl in gauss_c3 gauss_c2 gauss_c1 box_c5 gauss_c3_pffft; do cut -c1-400 $O/${TAG}_bench_$wl.json; tail -1 $O/${TAG}_bench_$wl.err; done
cut -c1-300 $O/${TAG}_bench_ref.json; wc -l $O/${TAG}_configs.jsonl $O/${TAG}_sweep_flag*.jsonl
