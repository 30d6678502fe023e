#!/bin/bash
# Builds copies of the library that differ only in like_lean.cu's compile flags, for A/B timing on the GPU box:
#   scripts/build_variants.sh name1 "flags1" name2 "flags2" ...   ->  _variants/lib_<name>.so  (WFPT_B200_LIB selects one)
set -e
cd "$(dirname "$0")/.."
python -m hddm_b200.build > /dev/null
mkdir -p _variants
ARCH="-gencode arch=compute_100a,code=sm_100a"
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  nvcc $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -I include -I hddm_b200/csrc $flags -Xptxas=-v -c hddm_b200/csrc/like_lean.cu -o /tmp/like_lean_$name.o 2> /tmp/ptxas_$name.txt
  objs=$(ls hddm_b200/csrc/_obj/*.o | grep -v like_lean)
  nvcc $ARCH -shared -o _variants/lib_$name.so $objs /tmp/like_lean_$name.o -cudart static
  echo "$name: $(grep -A2 'like_lean_kernel' /tmp/ptxas_$name.txt | grep -o 'Used [0-9]* registers\|[0-9]* bytes spill stores' | tr '\n' ' ')"
done
