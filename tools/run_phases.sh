set -x
python bench.py --workload fit5k --steps 40 --no-cpu-baseline --no-extras --no-parity > gpurun_out/r2c_fit5k.json 2>gpurun_out/r2c_fit5k.err; tail -2 gpurun_out/r2c_fit5k.err
python -c "
import json;d=json.load(open('gpurun_out/r2c_fit5k.json'));print(d['value'],d['stages']);print(d["roofline"]["ms_per_kernel"])"
timeout 120 python -m pytest tests/test_lbs_gpu.py tests/test_fit_gpu.py -m gpu -q -x 2>&1 | tail -2
UUO_EXTRA_NVCC_FLAGS=-DNP_TIMING python -m uuo_mocap_b200.build --force > /dev/null 2>&1
python tools/nn_phases.py 5000 prior mask hint
