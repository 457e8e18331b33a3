"""Summarise one kernel of an .ncu-rep (read with the ncu CLI here): headline metrics, stall
breakdown and the hottest SASS blocks.  usage: python tools/ncu_summary.py REPORT.ncu-rep [units]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
units_processed = float(sys.argv[2]) if len(sys.argv) > 2 else None
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
m = {h: (vals[i], units[i]) for i, h in enumerate(hdr)}
def g(k):
    return m.get(k, ('', ''))
print('kernel:', g('Kernel Name')[0], ' grid', g('launch__grid_size')[0], 'block', g('launch__block_size')[0])
keys = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__inst_executed.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__bytes_read.sum.per_second',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum', 'lts__t_sector_hit_rate.pct',
        'l1tex__t_bytes.sum', 'l1tex__t_sector_hit_rate.pct', 'smsp__inst_executed_op_local_ld.sum',
        'smsp__inst_executed_op_local_st.sum', 'smsp__inst_executed_op_shared_ld.sum', 'smsp__inst_executed_op_global_ld.sum']
for k in keys:
    v, u = g(k)
    if v != '':
        print('  %-62s %s %s' % (k, v, u))
def num(k):
    v, u = g(k)
    v = float(v.replace(',', ''))
    return v * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1, 'Tbyte': 1e12}.get(u, 1)
try:
    tr = num('dram__bytes_read.sum') + num('dram__bytes_write.sum')
    print('  dram traffic (read+write) per launch: %.4g bytes' % tr)
    if units_processed:
        print('  dram traffic per unit: %.1f bytes (%g units)' % (tr / units_processed, units_processed))
except Exception as e:
    print('  traffic: n/a', e)
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]; ci = {h: i for i, h in enumerate(hdr)}
data = []
for r in rows[2:]:
    if len(r) < 20: continue
    try:
        d = dict(addr=r[0], sass=r[1], samp=int(r[ci['# Samples']] or 0), inst=int(r[ci['Instructions Executed']] or 0),
                 thr=float(r[ci['Avg. Threads Executed']] or 0))
        for k in ('stall_long_sb', 'stall_short_sb', 'stall_wait', 'stall_branch_resolving', 'stall_no_inst', 'stall_selected',
                  'stall_not_selected', 'stall_math', 'stall_lg', 'stall_mio', 'stall_dispatch', 'stall_barrier'):
            d[k] = int(r[ci[k]] or 0)
        data.append(d)
    except Exception:
        pass
tot = sum(d['samp'] for d in data) or 1
toti = sum(d['inst'] for d in data) or 1
print('warp-stall samples: %d; stall reasons (%% of samples):' % tot)
print('  ' + '; '.join('%s %.1f' % (k[6:], 100.0 * sum(d[k] for d in data) / tot) for k in data[0] if k.startswith('stall_')))
blocks, cur = [], []
for i, d in enumerate(data):
    cur.append(i)
    if any(x in d['sass'] for x in ('BRA', 'EXIT', 'BSYNC', 'BREAK', 'RET', 'CALL')):
        blocks.append(cur); cur = []
if cur: blocks.append(cur)
print('hottest SASS basic blocks (share of samples, share of warp instructions, mean active threads, opcode mix):')
bl = []
for b in blocks:
    s = sum(data[i]['samp'] for i in b); n = sum(data[i]['inst'] for i in b); t = sum(data[i]['inst'] * data[i]['thr'] for i in b)
    bl.append((s, n, t, b))
for s, n, t, b in sorted(bl, key=lambda x: -x[0])[:14]:
    ops = {}
    for i in b:
        parts = data[i]['sass'].split()
        o = (parts[1] if parts[0].startswith('@') else parts[0]).split('.')[0]
        ops[o] = ops.get(o, 0) + 1
    top = ' '.join('%s:%d' % kv for kv in sorted(ops.items(), key=lambda x: -x[1])[:6])
    print('  @%s len %3d  samples %5.1f%%  inst %5.1f%%  threads %4.1f | %s' % (data[b[0]]['addr'][-5:], len(b), 100.0 * s / tot, 100.0 * n / toti, t / max(n, 1), top))
