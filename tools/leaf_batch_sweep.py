# prints kernel_ms for several leaf-batch values
import os, sys, subprocess, json
for lb in ("1","6","10","16"):
    env=dict(os.environ, MPENV_LEAF_BATCH=lb)
    out=subprocess.run([sys.executable,"bench.py","--steps","50","--warmup","5","--no-cpu-baseline"],capture_output=True,text=True,env=env)
    try:
        j=json.loads(out.stdout.strip().splitlines()[-1]); print(lb, j["value"], j["roofline"]["kernel_ms"])
    except Exception as e: print(lb,"ERR",out.stderr[-500:])
