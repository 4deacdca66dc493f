#!/bin/bash
# ncu --set full of the three dominant kernels of the default workload (fit100k, N=1):  gpurun -- bash tools/profile_top3.sh r2g
TAG=${1:-rX}
ARGS="--steps 4 --warmup 3 --no-cpu-baseline --no-extras --no-parity"
python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.txt 2>&1; tail -3 gpurun_out/${TAG}_pytest_gpu.txt
python __graft_entry__.py --smoke > gpurun_out/${TAG}_smoke.txt 2>&1; tail -1 gpurun_out/${TAG}_smoke.txt
python bench.py $ARGS > gpurun_out/${TAG}_plain2.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:umma_gemm_kernel|skin_tc_kernel|nn_presorted_kernel|mean_dist_kernel' --launch-skip 6 --launch-count 5 -f -o gpurun_out/${TAG}_top3 python bench.py $ARGS > gpurun_out/ncu3.log 2>&1
grep -c "==PROF==" gpurun_out/ncu3.log
