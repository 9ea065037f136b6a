#!/bin/bash
# usage (on the GPU box): tools/variants.sh [-t] "<EXTRA flags 1>" "<EXTRA flags 2>" ...   -- rebuilds the product
# library with each set of flags and prints the bench's step / kernel times (-t: run the lidar / combat parity tests too).
T=0; if [ "$1" = "-t" ]; then T=1; shift; fi
for v in "$@"; do
  (cd madrona-mp-env_b200 && rm -f libmpenv_b200.so && make libmpenv_b200.so EXTRA="$v" > /dev/null 2>&1)
  echo "== $v"
  grep -E "Compiling entry function '_ZN5mpenv(14k_lidar_packet|10k_los_pool)" -A4 madrona-mp-env_b200/build.log | grep -E "registers" | tr '\n' ' '; echo
  if [ $T = 1 ]; then timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "lidar_packets or combat or stress" 2>&1 | tail -2; fi
  timeout 300 python bench.py --no-cpu-baseline --steps 200 --warmup 20 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'])"
done
