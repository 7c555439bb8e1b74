#!/bin/bash
O=gpurun_out; mkdir -p $O; T=r2c
timeout 900 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${T}_pytest.log
timeout 600 python tools/diag/pass_time.py --variants "OLS2=0;OLS2=1" > $O/${T}_pass_c3.jsonl 2> $O/${T}_pass.err
BLUR_B200_LIB=$PWD/blur_algorithms_b200/libblur_b200_c2.so timeout 600 python tools/diag/pass_time.py --variants "OLS2=1" >> $O/${T}_pass_c3.jsonl 2>> $O/${T}_pass.err
timeout 300 python tools/diag/pass_time.py --rows 2160 --cols 3840 --sigma 25 --reps 30 --variants "OLS2=1" > $O/${T}_pass_c2.jsonl 2>> $O/${T}_pass.err
timeout 300 python tools/diag/pass_time.py --rows 1080 --cols 1920 --sigma 5 --reps 50 --variants "OLS2=1" > $O/${T}_pass_c1.jsonl 2>> $O/${T}_pass.err
tail -5 $O/${T}_pytest.log; cat $O/${T}_pass_c3.jsonl $O/${T}_pass_c2.jsonl $O/${T}_pass_c1.jsonl; tail -3 $O/${T}_pass.err
bash tools/gpu_ncu.sh ${T}
