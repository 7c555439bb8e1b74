O=gpurun_out
python tools/diag/p2p_bw.py > $O/r01g_p2p.txt 2>&1
for f in 1 0; do
BLUR_B200_FUSED=$f timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 --no-e2e > $O/r01g_bench_n2_fused$f.json 2> $O/r01g_bench_n2_fused$f.err
done
tail -8 $O/r01g_p2p.txt; grep -h '^{' $O/r01g_bench_n2_fused*.json | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['config'].get('exchange')[:20], d['ms_per_step'], d['roofline']['row_pass_ms'], d['roofline']['col_pass_plus_interleave_ms'])"
