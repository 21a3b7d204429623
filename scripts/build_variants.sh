#!/bin/bash
# A/B builds of the library that differ only in compile-time switches of one translation unit:
#   scripts/build_variants.sh emp v1 "-DB2OT_TC_EPI_SINGLE=1" v2 "-DB2OT_TC_EPI_SINGLE=1 -DB2OT_TC_POLY=1" ...
# -> pot_b200/libb2ot_<name>.so (load with B2OT_LIB=pot_b200/libb2ot_<name>.so).  The other objects come from the
# regular build (make -C pot_b200/csrc).
set -e
cd "$(dirname "$0")/../pot_b200/csrc"
unit=$1; shift
make -j4 >/dev/null
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xptxas -v"
others=$(ls *.o | grep -v "^${unit}\.o$" | grep -v "_var_")
while [ $# -gt 1 ]; do
  name=$1; defs=$2; shift 2
  ( nvcc $FLAGS $defs -c ${unit}.cu -o ${unit}_var_${name}.o 2> ${unit}_var_${name}.ptxas.log || (tail -30 ${unit}_var_${name}.ptxas.log; false)
    nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libb2ot_${name}.so ${unit}_var_${name}.o $others
    echo "built libb2ot_${name}.so ($defs)" ) &
done
wait
