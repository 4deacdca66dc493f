#!/bin/bash
# ncu --set full of the smaller kernels of the default workload (one launch each):  gpurun -- bash tools/profile_small.sh <tag>
TAG=${1:-rX}
ARGS="--steps 4 --warmup 3 --no-cpu-baseline --no-extras --no-parity"
python bench.py $ARGS > gpurun_out/${TAG}_plain3.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:fk_backward_kernel|fk_forward_kernel|rows_backward_kernel|rows_tc_bwd_kernel|rows_tc_fwd_kernel|adam_dev|temporal_kernel' --launch-skip 14 --launch-count 9 -f -o gpurun_out/${TAG}_small python bench.py $ARGS > gpurun_out/ncu4.log 2>&1
grep -c "==PROF==" gpurun_out/ncu4.log
