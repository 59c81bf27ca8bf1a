#!/bin/bash
# Builds experiment variants of the library: libmuscle_b200.so with infonce.cu compiled under
# -DMB_EXP=<n> for every n given, into muscle_b200/lib/exp/libmuscle_b200_<n>.so (git-ignored).
# MB_EXTRA_FLAGS=-DMB_PAIR_TIMING=1 adds the pair kernel's cycle counters (MUSCLE_PAIR_TIMING=1 prints them).
# On the GPU box: cp muscle_b200/lib/exp/libmuscle_b200_<n>.so muscle_b200/lib/libmuscle_b200.so
set -e
cd "$(dirname "$0")/.."
python -c "import muscle_b200.build as b; b.build()" > /dev/null
mkdir -p muscle_b200/lib/exp
for n in "$@"; do
  (
  nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Xcompiler -fPIC -DMB_EXP=$n $MB_EXTRA_FLAGS \
       -c muscle_b200/csrc/infonce.cu -o muscle_b200/lib/exp/infonce_$n.o
  objs=$(ls muscle_b200/lib/obj/*.o | grep -v "/infonce.o")
  nvcc -shared -o muscle_b200/lib/exp/libmuscle_b200_$n.so muscle_b200/lib/exp/infonce_$n.o $objs \
       -gencode arch=compute_100a,code=sm_100a
  ) &
done
wait
ls -la muscle_b200/lib/exp/*.so
