# usage: bash tools/diag/exchange_modes.sh N  — identity check + bench for each exchange mode at N GPUs
N=${1:-2}; O=gpurun_out; mkdir -p $O
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/diag/dist_check.py 2>&1 | grep -v "^\*\|^W\|NCCL" | tail -6
for mode in "push 2" "push 4" "push 8" "store 0"; do
set -- $mode
BLUR_B200_EXCHANGE=$1 BLUR_B200_CHUNKS=$2 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 --no-e2e 2> $O/xm_$1_$2.err | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('$mode', 'N=$N', 'ms/step', round(d['ms_per_step'],3), 'row', round(d['roofline']['row_pass_ms'],3), 'col', round(d['roofline']['col_pass_plus_interleave_ms'],3))"
tail -3 $O/xm_$1_$2.err | grep -i "error\|Traceback" 
done
