b() { python bench.py --no-cpu-baseline --steps 200 --warmup 20 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'])"; }
echo base; b
for w in 24 20; do echo "overlap los_warps=$w"; MPENV_OVERLAP=1 MPENV_OVL_LOS_WARPS=$w b; done
echo "overlap flat prio 24"; MPENV_OVERLAP=1 MPENV_OVL_LOS_WARPS=24 MPENV_OVL_FLAT=1 b
echo base; b
