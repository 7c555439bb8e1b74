N=${1:-2}; O=gpurun_out; mkdir -p $O
run() { env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --no-e2e 2> $O/pd.err | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('$*', 'N=$N', 'ms/step', round(d['ms_per_step'],3), 'row', round(d['roofline']['row_pass_ms'],3), 'col', round(d['roofline']['col_pass_plus_interleave_ms'],3))"; }
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/diag/dist_check.py 2>&1 | grep identical
for k in ${2:-4 8 12 16}; do run BLUR_B200_CHUNKS=$k; done
