# quick device-timed fit5k line: bash tools/quick_fit5k.sh <tag>
python bench.py --workload fit5k --steps 40 --no-cpu-baseline --no-extras --no-parity > gpurun_out/$1_fit5k.json 2>gpurun_out/$1_fit5k.err; tail -2 gpurun_out/$1_fit5k.err
python - <<PY
import json
d=json.load(open('gpurun_out/$1_fit5k.json'))
print(d['value'], {k:round(v['ms_per_iter'],4) for k,v in d['stages'].items()}, 'e2e', d['e2e']['value'])
print(d['roofline']['ms_per_kernel'])
PY
