#!/bin/bash
# On the GPU box: times the headline launch with every library under _variants/ (see build_variants.sh); "adaptive" adds those runs.
for lib in _variants/lib_*.so; do
  echo "== $lib"
  WFPT_B200_LIB=$PWD/$lib python scripts/time_fixed.py $1 2>&1 | tail -4
done
