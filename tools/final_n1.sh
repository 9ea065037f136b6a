#!/bin/bash
# usage (on the GPU box): tools/final_n1.sh <tag>   -- GPU test suite, the N = 1 bench lines, one ncu --set full capture of the six step kernels
tag=$1
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/${tag}_pytest_gpu.txt
python bench.py 2>/dev/null | tail -1 > gpurun_out/${tag}_bench_w8192_n1.json
python bench.py --steps 20 --warmup 5 2>/dev/null | tail -1 > gpurun_out/${tag}_bench_driver_steps20_warmup5.json
python bench.py --no-cpu-baseline --worlds 4096 2>/dev/null | tail -1 > gpurun_out/${tag}_bench_w4096_n1.json
python bench.py --no-cpu-baseline --workload stress --worlds 16384 2>/dev/null | tail -1 > gpurun_out/${tag}_bench_stress_w16384_n1.json
for f in w8192_n1 driver_steps20_warmup5 w4096_n1 stress_w16384_n1; do python -c "
import json,sys; d=json.load(open('gpurun_out/${tag}_bench_$f.json')); print('$f', d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['kernel_ms'])"; done
python tools/prof_step.py 8192 60 2 > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_(move|fire|world|los|obs|lidar)" -s 366 -c 6 -f -o gpurun_out/${tag}_step_full python tools/prof_step.py 8192 60 2 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
