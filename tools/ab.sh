#!/bin/bash
# Same-box A/B of two builds of the library (box-to-box clock differences are +-4 %):
#   cp uuo-mocap_b200/lib/libuuomocap.so uuo-mocap_b200/lib/libuuomocap_base.so   (before the change), rebuild, then
#   gpurun -- bash tools/ab.sh python tools/nearest_points_phases.py 100000 1
BASE=uuo-mocap_b200/lib/libuuomocap_base.so
for rep in 1 2; do
  echo "== base ($rep)"; UUO_MOCAP_LIB=$PWD/$BASE "$@" 2>&1 | grep -v "^$" | tail -${AB_TAIL:-6}
  echo "== new  ($rep)"; "$@" 2>&1 | grep -v "^$" | tail -${AB_TAIL:-6}
done
