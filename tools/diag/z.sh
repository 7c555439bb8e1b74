O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
b() { env "$@" python bench.py --steps 10 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('$*'.ljust(60), 'ms/step', round(d['ms_per_step'],3), 'row', round(d['roofline']['row_pass_ms'],3), 'col+il', round(d['roofline']['col_pass_plus_interleave_ms'],3))"; }
b X=0
b BLUR_B200_ROW_CTAS=4 BLUR_B200_ROW_CARVEOUT=86
b BLUR_B200_ROW_CTAS=4 BLUR_B200_ROW_CARVEOUT=72
b BLUR_B200_ROW_CTAS=4
b BLUR_B200_OLS_R0=8
b BLUR_B200_OLS_R0=16
python tools/bench_configs.py --steps 10 > $O/r01g_configs.jsonl 2> $O/r01g_configs.err; tail -2 $O/r01g_configs.err
python - <<'PY'
import json
for l in open('gpurun_out/r01g_configs.jsonl'):
    d=json.loads(l); print(d['config'][:70].ljust(70), 'ms', round(d['ms'],3), 'Mpx/s', round(d['megapixels_per_s']), 'frac', round(d['frac'],3))
PY
