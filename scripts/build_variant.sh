#!/bin/bash
# Tuning aid: primate_b200/variants_<name>.so = the current objects with slq_lanczos_f64.cu recompiled under extra -D flags.
#   bash scripts/build_variant.sh <name> -DSLQ_TMA_UF=5 -DSLQ_TMA_RF=1      (run the default build first)
set -e
name=$1; shift
B=primate_b200/csrc/build
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
  --expt-relaxed-constexpr -Iinclude -Iprimate_b200/csrc -DSLQ_BUILD "$@" -Xptxas -v -c primate_b200/csrc/slq_lanczos_f64.cu -o $B/var_$name.o > $B/var_$name.log 2>&1
objs=$(ls $B/slq_*.o | grep -v slq_lanczos_f64.o)
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC $objs $B/var_$name.o -o primate_b200/variants_$name.so
grep -A1 "k_stepA_tmaIdLi64ELi4ELb0" $B/var_$name.log | grep -o "Used [0-9]* registers\|[0-9]* bytes stack frame, [0-9]* bytes spill stores" | tr '\n' ' '; echo
