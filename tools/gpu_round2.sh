#!/bin/bash
# One GPU-box visit (round 2): gpu tests, smoke, bench for every workload + reference arm, ncu launch lists,
# ncu --set full of the FFT kernels and of the box kernel.
# usage (from the repo root, under gpurun): bash tools/gpu_round2.sh TAG [skip-tests]
TAG=${1:-r02}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
nproc >> $O/${TAG}_smi.txt
if [ -z "$2" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest.log
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> $O/${TAG}_smoke.log
fi
for wl in gauss_c3 gauss_c2 gauss_c1 box_c5; do
  timeout 600 python bench.py --workload $wl --steps 20 --warmup 3 > $O/${TAG}_bench_$wl.json 2> $O/${TAG}_bench_$wl.err; echo "bench rc=$?" >> $O/${TAG}_bench_$wl.err
done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/${TAG}_launches.csv \
   python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-verify > $O/${TAG}_ncu_launch.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/${TAG}_launches_box.csv \
   python bench.py --workload box_c5 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-verify > $O/${TAG}_ncu_launch_box.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'ols2?_(row|col)_kernel|interleave' -s 6 -c 3 -f -o $O/${TAG}_full \
   python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-verify > $O/${TAG}_ncu_full.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'box' -s 4 -c 2 -f -o $O/${TAG}_box_full \
   python bench.py --workload box_c5 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-verify > $O/${TAG}_ncu_box_full.log 2>&1
tail -3 $O/${TAG}_pytest.log $O/${TAG}_smoke.log 2>/dev/null
for wl in gauss_c3 gauss_c2 gauss_c1 box_c5; do cut -c1-1500 $O/${TAG}_bench_$wl.json; tail -2 $O/${TAG}_bench_$wl.err; done
cut -c1-600 $O/${TAG}_bench_ref.json
