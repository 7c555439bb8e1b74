#!/bin/bash
# Multi-GPU visit: the sharded-identity gpu test, then bench.py per exchange mode at every N <= NMAX.
# usage (under gpurun --gpus NMAX): bash tools/gpu_xmodes.sh TAG NMAX
TAG=${1:-xm}; NMAX=${2:-2}; O=gpurun_out; mkdir -p $O
nvidia-smi topo -m > $O/${TAG}_topo.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -k "sharded" > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest.log; tail -4 $O/${TAG}_pytest.log
for n in 2 4 8; do
  [ $n -le $NMAX ] || continue
  for mode in "store 0" "push 0"; do
    set -- $mode
    BLUR_B200_EXCHANGE=$1 BLUR_B200_CHUNKS=$2 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 \
      bench.py --gpus $n --steps 20 --warmup 3 --no-e2e > $O/${TAG}_n${n}_$1.json 2> $O/${TAG}_n${n}_$1.err
    grep -h '^{' $O/${TAG}_n${n}_$1.json | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('$mode', 'N=$n', 'ms/step', round(d['ms_per_step'],3), 'row', round(d['roofline']['row_pass_ms'],3), 'col', round(d['roofline']['col_pass_plus_interleave_ms'],3), 'verified', d['verified']['ok'])"
    grep -i "error\|Traceback" $O/${TAG}_n${n}_$1.err | head -3
  done
done
# FastBoxBlur batch dealt over the ranks (BASELINE config 5): image i -> rank i mod N
for n in 1 2 4 8; do
  [ $n -le $NMAX ] || continue
  if [ $n -eq 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29513"; fi
  timeout 400 $L bench.py --workload box_c5 --gpus $n --steps 20 --warmup 3 --no-cpu-baseline > $O/${TAG}_box_n${n}.json 2> $O/${TAG}_box_n${n}.err
  grep -h '^{' $O/${TAG}_box_n${n}.json | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('box_c5 N=$n', 'ms/step', round(d['ms_per_step'],3), 'Mpx/s', round(d['value']), 'e2e Mpx/s', round(d['e2e']['value']), 'verified', d['verified']['ok'])"
done
# occupancy variants of the two FFT kernels (tools/build_variant.sh), single GPU
for v in r2c2; do
  [ -f blur_algorithms_b200/variants/libblur_b200_$v.so ] || continue
  echo "variant $v:"; BLUR_B200_LIB=$PWD/blur_algorithms_b200/variants/libblur_b200_$v.so timeout 300 python tools/diag/pass_time.py --variants "OLS2=1" 2>&1 | tail -1
done
echo "stock:"; timeout 300 python tools/diag/pass_time.py --variants "OLS2=1" 2>&1 | tail -1
