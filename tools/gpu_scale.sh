#!/bin/bash
# Multi-GPU visit: bit-identity check of the sharded blur vs 1 GPU, then bench.py at N GPUs.
# usage (under gpurun --gpus N): bash tools/gpu_scale.sh TAG N
TAG=${1:-scale}; N=${2:-2}; O=gpurun_out; mkdir -p $O
nvidia-smi topo -m > $O/${TAG}_topo.txt 2>&1
for n in $(seq 2 $N); do
  case $n in 2|4|8) ;; *) continue;; esac
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
     tools/diag/dist_check.py > $O/${TAG}_check_n$n.log 2>&1; echo "check n=$n rc=$?" >> $O/${TAG}_check_n$n.log
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 \
     bench.py --gpus $n --steps 20 --warmup 3 > $O/${TAG}_bench_n$n.json 2> $O/${TAG}_bench_n$n.err; echo "bench n=$n rc=$?" >> $O/${TAG}_bench_n$n.err
  tail -6 $O/${TAG}_check_n$n.log; grep -h '^{' $O/${TAG}_bench_n$n.json; tail -2 $O/${TAG}_bench_n$n.err
done
