#!/bin/bash
# first GPU visit of round 2: parity of the two-factor kernels + variant timing
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/r2a_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r2a_pytest.log
timeout 600 python tools/diag/pass_time.py > $O/r2a_pass_c3.jsonl 2> $O/r2a_pass_c3.err
timeout 300 python tools/diag/pass_time.py --rows 2160 --cols 3840 --sigma 25 --reps 30 --variants "OLS2=0;OLS2=1" > $O/r2a_pass_c2.jsonl 2>> $O/r2a_pass_c3.err
timeout 300 python tools/diag/pass_time.py --rows 1080 --cols 1920 --sigma 5 --reps 50 --variants "OLS2=0;OLS2=1" > $O/r2a_pass_c1.jsonl 2>> $O/r2a_pass_c3.err
tail -5 $O/r2a_pytest.log; cat $O/r2a_pass_c3.jsonl $O/r2a_pass_c2.jsonl $O/r2a_pass_c1.jsonl; tail -3 $O/r2a_pass_c3.err
