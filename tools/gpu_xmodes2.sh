#!/bin/bash
# usage (under gpurun --gpus NMAX): bash tools/gpu_xmodes2.sh TAG "N list" "mode list"   mode = exchange[:stages]
TAG=${1:-xm}; NS=${2:-2}; MODES=${3:-"staged:4 staged:8 staged:2 store push"}; O=gpurun_out; mkdir -p $O
for n in $NS; do
  for m in $MODES; do
    ex=${m%%:*}; st=${m##*:}; [ "$st" = "$m" ] && st=0
    BLUR_B200_EXCHANGE=$ex BLUR_B200_STAGES=$st timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 \
      bench.py --gpus $n --steps 20 --warmup 3 --no-e2e > $O/${TAG}_n${n}_${ex}${st}.json 2> $O/${TAG}_n${n}_${ex}${st}.err
    grep -h '^{' $O/${TAG}_n${n}_${ex}${st}.json | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('$m', 'N=$n', 'ms/step', round(d['ms_per_step'],3), 'row', round(d['roofline']['row_pass_ms'],3), 'col', round(d['roofline']['col_pass_plus_interleave_ms'],3), 'verified', d['verified']['ok'])"
    grep -i "error\|Traceback" $O/${TAG}_n${n}_${ex}${st}.err | head -3
  done
done
