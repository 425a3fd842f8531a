"""Summarise an ncu raw+source CSV pair: python tools/ncu_summary.py raw.csv src.csv"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ('gpu__time_duration.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum', 'launch__registers_per_thread',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ldgsts.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum',
        'sm__warps_active.avg.per_cycle_active', 'launch__grid_size', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum')
for h, u, v in zip(hdr, units, vals):
    if h in want or ('pcsamp_warps_issue_stalled' in h and 'not_issued' not in h):
        print("%-75s %-10s %s" % (h, u, v))
if len(sys.argv) > 2:
    rows = list(csv.reader(open(sys.argv[2])))
    hdr = rows[1]; data = rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    by_op = collections.Counter(); tot = 0; recs = []
    for r in data:
        if len(r) < len(hdr): continue
        try: s = int(r[ix['# Samples']])
        except Exception: continue
        src = r[ix['Source']].strip(); toks = src.split()
        op = toks[1] if toks and toks[0].startswith('@') else (toks[0] if toks else '?')
        by_op[op.split('.')[0]] += s; tot += s
        recs.append((s, r[ix['Address']][-5:], src[:70]))
    print("total samples", tot); print(by_op.most_common(14))
    for x in sorted(recs, reverse=True)[:16]: print(x)
