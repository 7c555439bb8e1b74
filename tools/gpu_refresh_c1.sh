#!/bin/bash
# Short visit after the direct kernel went in: gpu tests, smoke, the C1 bench line, the other configs, C1 launch list.
TAG=${1:-r02y}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> $O/${TAG}_smoke.log
timeout 300 python bench.py --workload gauss_c1 --steps 30 --warmup 3 > $O/${TAG}_bench_gauss_c1.json 2> $O/${TAG}_bench_gauss_c1.err
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_gauss_c3.json 2> $O/${TAG}_bench_gauss_c3.err
timeout 300 python tools/bench_configs.py > $O/${TAG}_configs.jsonl 2> $O/${TAG}_configs.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $O/${TAG}_launches_c1.csv \
   python bench.py --workload gauss_c1 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-verify > $O/${TAG}_ncu_launch_c1.log 2>&1
tail -3 $O/${TAG}_pytest.log $O/${TAG}_smoke.log; cut -c1-500 $O/${TAG}_bench_gauss_c1.json; cut -c1-200 $O/${TAG}_bench_gauss_c3.json; head -2 $O/${TAG}_configs.jsonl | cut -c1-160
