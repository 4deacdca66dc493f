# default bench at N=1 without the slow legs: bash tools/quick_default.sh <tag>
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras $2 > gpurun_out/$1_q.json 2>gpurun_out/$1_q.err; tail -2 gpurun_out/$1_q.err
python - <<PY
import json
d=json.load(open('gpurun_out/$1_q.json'))
print('value %.3f M  e2e %.3f M' % (d['value']/1e6, d['e2e']['value']/1e6), d['e2e']['seconds_per_rank'], d['e2e']['phase_ms_rank0'])
print({k:round(v['ms_per_iter'],3) for k,v in d['stages'].items()}, d['clocks'])
PY
