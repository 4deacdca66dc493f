#!/bin/bash
# One profiling pass on the GPU box for profiles/: bench lines, the ncu launch list and --set full captures of the
# top kernels of the DEFAULT workload (fit100k).   gpurun -- bash tools/profile_round.sh r2g ; then
# (FULL=1 adds a --set full capture of every kernel of the iteration; tools/profile_top3.sh has the dominant ones)
# python tools/ncu_summary.py r2g gpurun_out/r2g_launches.csv gpurun_out/r2g_full.ncu-rep   here
TAG=${1:-rX}
mkdir -p gpurun_out
ARGS="--steps 4 --warmup 3 --no-cpu-baseline --no-extras --no-parity"
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || tail -5 gpurun_out/${TAG}_bench.err
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/${TAG}_reference.json 2> gpurun_out/${TAG}_reference.err || tail -5 gpurun_out/${TAG}_reference.err
python bench.py $ARGS > gpurun_out/${TAG}_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py $ARGS > gpurun_out/ncu1.log 2>&1
[ -n "$FULL" ] && timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:umma_gemm_kernel|skin_tc_kernel|nn_presorted_kernel|rows_backward|rows_forward|fk_backward|fk_forward|adam_dev' --launch-skip 18 --launch-count 10 -f -o gpurun_out/${TAG}_full python bench.py $ARGS > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu1.log gpurun_out/ncu2.log
cut -c1-400 gpurun_out/${TAG}_bench.json; cut -c1-300 gpurun_out/${TAG}_reference.json
