#!/bin/bash
# Scaling lines of the fitting workloads on N GPUs of one box:  gpurun --gpus N -- bash tools/scale_run.sh N [steps]
N=${1:-2}; K=${2:-20}
mkdir -p gpurun_out
run() {  # name, extra args
  if [ "$N" = "1" ]; then python bench.py --gpus 1 "${@:2}" > gpurun_out/scale_$1_n$N.json 2> gpurun_out/scale_$1_n$N.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
       bench.py --gpus $N "${@:2}" > gpurun_out/scale_$1_n$N.json 2> gpurun_out/scale_$1_n$N.err; fi
  tail -c 300 gpurun_out/scale_$1_n$N.err | tail -2; cut -c1-200 gpurun_out/scale_$1_n$N.json
}
run fit100k --workload fit100k --steps $K --no-cpu-baseline
run mht --workload mht --steps $K --no-cpu-baseline
run lbs --steps 100 --no-cpu-baseline
