"""Join an ncu source page (SASS view, --csv) with nvdisasm -g line info: warp instructions, active lanes and stall
samples per source line.  usage: python tools/ncu_lines2.py SOURCE.csv DISASM.txt MANGLED_FUNCTION [top]"""
import csv, re, sys
src, dis, fun = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
line_of = {}
cur, inside = None, False
for l in open(dis):
    if l.startswith('.text.'):
        inside = l.strip().rstrip(':') == '.text.' + fun
        continue
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'\s+/\*([0-9a-f]+)\*/', l)
    if m: line_of[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(src)))
hdr = rows[1]; ci = {h: i for i, h in enumerate(hdr)}
agg = {}
base = None
tot_i = tot_s = 0
for r in rows[2:]:
    if len(r) < 20: continue
    try:
        a = int(r[0], 16)
    except ValueError:
        continue
    if base is None: base = a
    inst = int(r[ci['Instructions Executed']] or 0); thr = int(r[ci['Thread Instructions Executed']] or 0)
    samp = int(r[ci['# Samples']] or 0)
    k = line_of.get(a - base, ('?', 0))
    e = agg.setdefault(k, [0, 0, 0, 0])
    e[0] += inst; e[1] += thr; e[2] += samp; e[3] += 1
    tot_i += inst; tot_s += samp
print('total warp instructions %.4g, samples %d' % (tot_i, tot_s))
byfile = {}
for (f, ln), e in agg.items():
    b = byfile.setdefault(f, [0, 0, 0])
    b[0] += e[0]; b[1] += e[1]; b[2] += e[2]
for f, b in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print('%-28s inst %5.1f%%  lanes %4.1f  samples %5.1f%%' % (f, 100.0 * b[0] / tot_i, b[1] / max(b[0], 1), 100.0 * b[2] / tot_s))
print()
for (f, ln), e in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print('%-24s:%-5d sass %3d  inst %5.2f%%  lanes %4.1f  samples %5.2f%%' % (f, ln, e[3], 100.0 * e[0] / tot_i, e[1] / max(e[0], 1), 100.0 * e[2] / tot_s))
