#!/bin/bash
# One profiling pass on the GPU box: bench lines, ncu launch lists and --set full captures for profiles/.
#   gpurun -- bash tools/profile_round.sh r1f ; then python tools/ncu_summary.py ... here
TAG=${1:-rX}
mkdir -p gpurun_out
set -x
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || tail -5 gpurun_out/${TAG}_bench.err
python bench.py --workload fit5k > gpurun_out/${TAG}_fit5k.json 2> gpurun_out/${TAG}_fit5k.err || tail -5 gpurun_out/${TAG}_fit5k.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_fit_launches.csv python bench.py --workload fit5k --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:umma_gemm|skin_tc|skin_bwdA|fk_' --launch-skip 21 --launch-count 7 -f -o gpurun_out/${TAG}_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu3.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:nn_presorted|rows_|chamfer_rows|adam' --launch-skip 21 --launch-count 7 -f -o gpurun_out/${TAG}_fit_full python bench.py --workload fit5k --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu4.log 2>&1
tail -2 gpurun_out/ncu3.log gpurun_out/ncu4.log
cat gpurun_out/${TAG}_bench.json | cut -c1-300
