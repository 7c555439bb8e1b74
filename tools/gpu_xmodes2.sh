#!/bin/bash
# usage (under gpurun --gpus NMAX): bash tools/gpu_xmodes2.sh TAG "N list" "mode list"
#   mode = exchange[:stages[:row_ctas[:col_ctas]]]   (row/col CTAs per SM: sstore only)
TAG=${1:-xm}; NS=${2:-2}; MODES=${3:-"staged:4 staged:8 staged:2 store push"}; O=gpurun_out; mkdir -p $O
for n in $NS; do
  for m in $MODES; do
    IFS=: read ex st rc cc <<< "$m"; st=${st:-0}; rc=${rc:-2}; cc=${cc:-1}
    tag=${TAG}_n${n}_${ex}${st}_r${rc}c${cc}
    BLUR_B200_EXCHANGE=$ex BLUR_B200_STAGES=$st BLUR_B200_SSTORE_ROW_CTAS=$rc BLUR_B200_SSTORE_COL_CTAS=$cc timeout 300 \
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 \
      bench.py --gpus $n --steps 20 --warmup 3 --no-e2e > $O/$tag.json 2> $O/$tag.err
    grep -h '^{' $O/$tag.json | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('$m', 'N=$n', 'ms/step', round(d['ms_per_step'],3), 'row', round(d['roofline']['row_pass_ms'],3), 'col', round(d['roofline']['col_pass_plus_interleave_ms'],3), 'verified', d['verified']['ok'])"
    grep -i "error\|Traceback" $O/$tag.err | head -3
  done
done
