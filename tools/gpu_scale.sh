#!/bin/bash
# Multi-GPU visit: bit-identity check of the sharded blur vs 1 GPU, then bench.py at N GPUs (the contract sharding with the
# halo line beside it), then the image-sharded FastBoxBlur batch (BASELINE config 5) at the same N.
# usage (under gpurun --gpus N): bash tools/gpu_scale.sh TAG N [nocheck]
TAG=${1:-scale}; N=${2:-2}; CHECK=${3:-check}; O=gpurun_out; mkdir -p $O
nvidia-smi topo -m > $O/${TAG}_topo.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 8 4 2; do
  [ $n -le $N ] || continue
  if [ $CHECK = check ]; then
    timeout 300 $TR --nproc-per-node $n --master-port 29511 tools/diag/dist_check.py > $O/${TAG}_check_n$n.log 2>&1
    echo "check n=$n rc=$?" >> $O/${TAG}_check_n$n.log; grep -h "identical\|rc=" $O/${TAG}_check_n$n.log | tail -6
  fi
  timeout 400 $TR --nproc-per-node $n --master-port 29512 bench.py --gpus $n --steps 20 --warmup 3 \
     > $O/${TAG}_bench_n$n.json 2> $O/${TAG}_bench_n$n.err; echo "bench n=$n rc=$?" >> $O/${TAG}_bench_n$n.err
  timeout 300 $TR --nproc-per-node $n --master-port 29513 bench.py --gpus $n --steps 20 --warmup 3 --workload box_c5 --no-e2e \
     > $O/${TAG}_box_n$n.json 2> $O/${TAG}_box_n$n.err; echo "box n=$n rc=$?" >> $O/${TAG}_box_n$n.err
  python - <<PY
import json
for f in ("$O/${TAG}_bench_n$n.json", "$O/${TAG}_box_n$n.json"):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print($n, d["config"]["workload"], round(d["ms_per_step"], 4), d.get("run", {}).get("exchange_mode"), d.get("run", {}).get("exchange_trial_ms"),
              "halo", (d.get("halo") or {}).get("ms_per_step"), ((d.get("halo") or {}).get("verified") or {}).get("ok"), "verified", (d.get("verified") or {}).get("ok"),
              "e2e", (d.get("e2e") or {}).get("value"))
    except Exception as e:
        print($n, f, "no line:", e)
PY
  tail -2 $O/${TAG}_bench_n$n.err | cut -c1-300
done
