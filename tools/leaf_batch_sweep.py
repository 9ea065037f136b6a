# prints kernel_ms for several leaf-batch values
import os, sys, subprocess, json
for lb, rb in (("1","1"),("10","1"),("1","4"),("10","4"),("10","8")):
    env=dict(os.environ, MPENV_LEAF_BATCH=lb, MPENV_REFILL_BATCH=rb)
    out=subprocess.run([sys.executable,"bench.py","--steps","50","--warmup","5","--no-cpu-baseline"],capture_output=True,text=True,env=env)
    try:
        j=json.loads(out.stdout.strip().splitlines()[-1]); print(lb, rb, j["value"], j["roofline"]["kernel_ms"])
    except Exception as e: print(lb,"ERR",out.stderr[-500:])
