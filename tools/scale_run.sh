#!/bin/bash
# The driver's scaling run of the default workload (fit100k) on N GPUs of one box:  gpurun --gpus N -- bash tools/scale_run.sh N [steps] [tag]
N=${1:-2}; K=${2:-20}; TAG=${3:-scale}
mkdir -p gpurun_out
if [ "$N" = "1" ]; then python bench.py --gpus 1 --steps $K --warmup 5 > gpurun_out/${TAG}_n$N.json 2> gpurun_out/${TAG}_n$N.err
else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
     bench.py --gpus $N --steps $K --warmup 5 > gpurun_out/${TAG}_n$N.json 2> gpurun_out/${TAG}_n$N.err; fi
echo rc=$?; tail -c 400 gpurun_out/${TAG}_n$N.err | tail -3
python - <<PY
import json
d=json.load(open('gpurun_out/${TAG}_n$N.json'))
print('N=$N value %.3f M  ms/step %.3f  e2e %.3f M (%s)' % (d['value']/1e6, d['ms_per_step'], d['e2e']['value']/1e6, d['e2e'].get('seconds_per_rank')))
print(d['run'], {k:round(v['ms_per_iter'],3) for k,v in d['stages'].items()}, d['clocks'])
print(d['roofline']['ms_per_kernel'])
PY
