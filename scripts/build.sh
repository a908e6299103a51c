#!/bin/bash
# Builds aha_b200/lib/libaha_b200.so for sm_100a (cross-compiles without a GPU).
set -e
cd "$(dirname "$0")/../aha_b200/csrc"
mkdir -p ../lib ../../build
OBJ=../../build
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC"
pids=()
for f in *.cu *.cpp; do
  o="$OBJ/${f%.*}.o"
  if [ ! -f "$o" ] || [ "$f" -nt "$o" ] || [ -n "$(find . -maxdepth 1 \( -name '*.cuh' -o -name '*.h' \) -newer "$o" 2>/dev/null)" ] || [ ../../include/aha_b200.h -nt "$o" ]; then
    nvcc $FLAGS -I../../include -c "$f" -o "$o" &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
nvcc -shared -o ../lib/libaha_b200.so $OBJ/*.o
echo "built aha_b200/lib/libaha_b200.so"
