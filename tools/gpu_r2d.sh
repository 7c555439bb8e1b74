#!/bin/bash
# L2 / DRAM traffic experiments on the two-factor column kernel
O=gpurun_out; mkdir -p $O; T=r2d
M="dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct"
run() {  # name, env..., EXTRA args via $EXTRA
  name=$1; shift
  env "$@" timeout 300 ncu --metrics $M --clock-control none -k regex:'ols2_col' -s 2 -c 1 --csv --log-file $O/${T}_$name.csv \
      python tools/diag/pass_time.py --reps 1 --variants "OLS2=1" $EXTRA > $O/${T}_$name.log 2>&1
  echo "== $name"; grep -E "ols2_col" $O/${T}_$name.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}' | tr -d '"'
}
EXTRA=""
run base X=1
run fetch32 BLUR_B200_L2_FETCH=32
run fetch128 BLUR_B200_L2_FETCH=128
run hint128 BLUR_B200_LIB=$PWD/blur_algorithms_b200/libblur_b200_h1.so
run evictlast BLUR_B200_LIB=$PWD/blur_algorithms_b200/libblur_b200_h2.so
EXTRA="--sigma 5"
run sigma5 X=1
EXTRA=""
for v in "X=1" "BLUR_B200_ROW_BULK=0" "BLUR_B200_L2_FETCH=32" "BLUR_B200_L2_FETCH=128" "BLUR_B200_LIB=$PWD/blur_algorithms_b200/libblur_b200_h1.so" "BLUR_B200_LIB=$PWD/blur_algorithms_b200/libblur_b200_h2.so"; do
  echo "-- $v"; env $v timeout 300 python tools/diag/pass_time.py --variants "OLS2=1" 2>&1 | tail -1
done
timeout 600 python -m pytest tests -m gpu -x -q -k "not c3_full and not c2_c3" > $O/${T}_pytest.log 2>&1; tail -3 $O/${T}_pytest.log
