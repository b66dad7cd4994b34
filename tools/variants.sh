#!/bin/bash
# Build experimental variants of libspraygpu.so (same sources, different -D knobs) next to the default
# build, for A/B timing in one gpurun call:  SPRAYGPU_LIB=spray_b200/variants/libspraygpu_<name>.so python bench.py ...
# usage: tools/variants.sh name1 "-DFLAG=1" name2 "-DOTHER=2" ...
set -e
cd "$(dirname "$0")/../spray_b200/csrc"
mkdir -p ../variants
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  rm -f *.o
  make -s OUT=../variants/libspraygpu_$name.so EXTRA="$flags" > /dev/null
  grep -A2 "render_tile_kernelILb0" trace.ptxas.log | grep -E "registers|spill" | tr '\n' ' ' | sed "s/^/$name: /"; echo
done
rm -f *.o
make -s > /dev/null
