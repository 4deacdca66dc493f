#!/bin/bash
# ncu --set full of the mean-distance kernel at F = 100 000 (one launch):  gpurun -- bash tools/profile_meandist.sh <tag>
TAG=${1:-rX}
python tools/nearest_points_phases.py 100000 1 > gpurun_out/${TAG}_np_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:mean_dist_kernel' --launch-skip 5 --launch-count 1 -f -o gpurun_out/${TAG}_meandist python tools/nearest_points_phases.py 100000 1 > gpurun_out/ncu5.log 2>&1
grep -c "==PROF==" gpurun_out/ncu5.log; tail -12 gpurun_out/${TAG}_np_plain.log | cut -c1-200
