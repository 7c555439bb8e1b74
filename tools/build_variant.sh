#!/bin/bash
# Builds an experiment variant of the library: tools/build_variant.sh NAME "-DFLAG=..." [file.cu ...]
# Recompiles the named sources (default ols2.cu) with the extra flags and links them with the stock objects
# into blur_algorithms_b200/variants/libblur_b200_NAME.so (load it with BLUR_B200_LIB=...).
set -e
NAME=$1; FLAGS=$2; shift 2 || true
SRCS=${@:-ols2.cu}
cd "$(dirname "$0")/../blur_algorithms_b200/csrc"
unset CC CXX
make -s -j8 > /dev/null
mkdir -p build/var_$NAME ../variants
OBJS=""
for o in build/*.o; do
  b=$(basename $o .o)
  if echo " $SRCS " | grep -q " $b.cu "; then
    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -Xptxas -v $FLAGS \
      -c $b.cu -o build/var_$NAME/$b.o 2> build/var_$NAME/$b.ptxas.log || { cat build/var_$NAME/$b.ptxas.log; exit 1; }
    OBJS="$OBJS build/var_$NAME/$b.o"
  else
    OBJS="$OBJS $o"
  fi
done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libblur_b200_$NAME.so $OBJS
grep -A1 "ols2_row_kernelILi32ELb0ELb1E\|ols2_col_kernel" build/var_$NAME/ols2.ptxas.log 2>/dev/null | grep -E "registers|spill" | head -4
