#!/usr/bin/env python
"""profiles/ helper: turns `ncu -i X.ncu-rep --page raw --csv` output into the per-round
summary table, the raw copy and profiles/ncu_traffic.json (what bench.py's roofline.traffic reads).
usage: ncu -i rep --page raw --csv > raw.csv; python tools/ncu_summary.py raw.csv r2_v1 [worlds]
ncu_traffic.json is stamped with a digest of the kernel sources it was captured on (bench.py
refuses to report traffic from another build)."""
import csv, json, os, sys
raw, tag = sys.argv[1], sys.argv[2]
worlds = int(sys.argv[3]) if len(sys.argv) > 3 else 8192
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import csrc_digest
rows = list(csv.reader(open(raw)))
h, units, data = rows[0], rows[1], rows[2:]
ix = {k: i for i, k in enumerate(h)}
keep = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'dram__bytes.sum.per_second', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_st.sum',
        'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic']
out = [['Kernel Name', ''] + [r[ix['Kernel Name']][:26] for r in data]]
for k in keep:
    if k in ix:
        out.append([k, units[ix[k]]] + [r[ix[k]] for r in data])
csv.writer(open(os.path.join(ROOT, 'profiles', f'{tag}_ncu_full_summary.csv'), 'w')).writerows(out)
open(os.path.join(ROOT, 'profiles', f'{tag}_ncu_full_raw.csv'), 'w').write(open(raw).read())
mult = {'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1, 'Gbyte': 1e9}
tr = {}
for r in data:
    n = r[ix['Kernel Name']]
    key = 'k_los' if 'k_los' in n else 'k_lidar' if 'k_lidar' in n else 'k_fire' if 'k_fire' in n else n.split('(')[0].replace('void ', '').split('<')[0]
    tr[key] = int(sum(float(r[ix[m]]) * mult[units[ix[m]]] for m in ('dram__bytes_read.sum', 'dram__bytes_write.sum')))
json.dump({"source": f"profiles/{tag}_ncu_full_summary.csv (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch, {worlds} worlds)",
           "worlds": worlds, "csrc_sha16": csrc_digest(), "bytes_per_launch": tr}, open(os.path.join(ROOT, 'profiles', 'ncu_traffic.json'), 'w'), indent=1)
for row in out[:10]:
    print(row)
