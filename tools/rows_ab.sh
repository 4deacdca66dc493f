# default bench with the marker-stage rows path on the tensor cores vs the shared-memory kernels: bash tools/rows_ab.sh <tag>
for k in tc smem; do
  if [ $k = smem ]; then export UUO_ROWS_SMEM=1; else unset UUO_ROWS_SMEM; fi
  python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/$1_$k.json 2>gpurun_out/$1_$k.err; tail -2 gpurun_out/$1_$k.err
  python - <<PY
import json
d=json.load(open('gpurun_out/$1_$k.json'))
print('$k value %.3f M  e2e %.3f M' % (d['value']/1e6, d['e2e']['value']/1e6), {k:round(v['ms_per_iter'],3) for k,v in d['stages'].items()}, d['clocks']['sm_mhz'])
print(d['roofline']['ms_per_kernel']['marker_iteration'])
print(d.get('parity'))
PY
done
