ARGS="--workload fit5k --steps 4 --warmup 3 --no-cpu-baseline --no-extras --no-parity"
python bench.py $ARGS > gpurun_out/g_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:umma_gemm --launch-skip 4 --launch-count 1 -f -o gpurun_out/$1_gemm python bench.py $ARGS > gpurun_out/ncu_g.log 2>&1
grep -c PROF gpurun_out/ncu_g.log
