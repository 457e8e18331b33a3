"""Attribute ncu warp-stall samples / instructions of one kernel to source lines using nvdisasm line
info of the object that holds it.  usage: python tools/ncu_lines.py REPORT.ncu-rep OBJECT.o KERNEL_SUBSTR"""
import csv, io, re, subprocess, sys, os, tempfile, glob, collections
rep, obj, kern = sys.argv[1], sys.argv[2], sys.argv[3]
d = tempfile.mkdtemp()
subprocess.run(['cuobjdump', '-xelf', 'all', os.path.abspath(obj)], cwd=d, capture_output=True)
cubin = glob.glob(os.path.join(d, '*.cubin'))[0]
txt = subprocess.run(['nvdisasm', '--print-line-info', cubin], capture_output=True, text=True).stdout
# locate the kernel's text section
m = re.search(r'\.section\s+\.text\.(\S*%s\S*),' % re.escape(kern), txt)
start = m.start()
end = txt.find('.section', start + 10)
sec = txt[start:end if end > 0 else None]
addr2line = {}
cur = None
for line in sec.splitlines():
    f = re.match(r'\s*//## File "([^"]+)", line (\d+)', line)
    if f:
        cur = (os.path.basename(f.group(1)), int(f.group(2)))
        continue
    a = re.match(r'\s*/\*([0-9a-f]{4,})\*/', line)
    if a and cur:
        addr2line[int(a.group(1), 16)] = cur
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]; ci = {h: i for i, h in enumerate(hdr)}
agg = collections.defaultdict(lambda: [0, 0, 0.0])
base = None
tot = toti = 0
for r in rows[2:]:
    if len(r) < 20: continue
    try:
        addr = int(r[0], 16)
    except Exception:
        continue
    if base is None: base = addr
    key = addr2line.get(addr - base, ('?', 0))
    s = int(r[ci['# Samples']] or 0); n = int(r[ci['Instructions Executed']] or 0); t = float(r[ci['Avg. Threads Executed']] or 0)
    a = agg[key]; a[0] += s; a[1] += n; a[2] += n * t
    tot += s; toti += n
srcs = {}
def text(f, l):
    if f not in srcs:
        p = [x for x in glob.glob(os.path.join(os.path.dirname(os.path.abspath(obj)), '..', '*')) if os.path.basename(x) == f]
        srcs[f] = open(p[0]).read().splitlines() if p else []
    return srcs[f][l - 1].strip()[:90] if 0 < l <= len(srcs[f]) else ''
byfile = collections.defaultdict(lambda: [0, 0])
for (f, l), a in agg.items():
    byfile[f][0] += a[0]; byfile[f][1] += a[1]
print('by file:', {f: '%.1f%% samples %.1f%% inst' % (100 * v[0] / tot, 100 * v[1] / toti) for f, v in byfile.items()})
for (f, l), a in sorted(agg.items(), key=lambda x: -x[1][0])[:int(sys.argv[4]) if len(sys.argv) > 4 else 40]:
    print('%-18s %4d  samples %5.1f%%  inst %5.1f%%  thr %4.1f | %s' % (f, l, 100 * a[0] / tot, 100 * a[1] / toti, a[2] / max(a[1], 1), text(f, l)))
