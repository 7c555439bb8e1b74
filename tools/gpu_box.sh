#!/bin/bash
cd /root/repo
O=gpurun_out; mkdir -p $O; T=r2f
timeout 600 python -m pytest tests -m gpu -x -q -k "box" > $O/${T}_pytest.log 2>&1; tail -15 $O/${T}_pytest.log
for v in 1 0; do
  BLUR_B200_BOX_PACKED=$v timeout 300 python bench.py --workload box_c5 --steps 10 --warmup 3 > $O/${T}_box_$v.json 2> $O/${T}_box_$v.err; tail -1 $O/${T}_box_$v.json | cut -c1-900; tail -2 $O/${T}_box_$v.err
done
