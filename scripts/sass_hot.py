"""Join an ncu SASS source page (csv) with nvdisasm --print-line-info to rank source lines by executed instructions,
stall samples and local-memory traffic.  usage: sass_hot.py <src.csv> <dis.txt> [top] [substring of the function's .text section name]"""
import csv, re, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
head = rows[1]
ci = {n: i for i, n in enumerate(head)}
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
fn = sys.argv[4] if len(sys.argv) > 4 else 'solve_kernel'
# address -> line
line_of = {}
cur = None
infunc = False
for l in open(sys.argv[2]):
    if '.text.' in l: infunc = fn in l
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2)))
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);', l)
    if m and infunc: line_of.setdefault(int(m.group(1), 16), cur)
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
base = None
for r in rows[2:]:
    try:
        a = int(r[ci['Address']], 16)
    except Exception:
        continue
    if base is None: base = a
    ln = line_of.get(a - base, ('?', 0))
    ie = int(r[ci['Instructions Executed']] or 0)
    smp = int(r[ci['# Samples']] or 0)
    loc = int(r[ci['L2 Theoretical Sectors Local']] or 0)
    src = r[ci['Source']]
    ag = agg[ln]
    ag[0] += ie; ag[1] += smp; ag[2] += loc
    if 'STL' in src or 'LDL' in src: ag[3] += ie
tot = [sum(v[i] for v in agg.values()) for i in range(4)]
print('total inst %.3g samples %d local sectors %.3g local inst %.3g' % (tot[0], tot[1], tot[2], tot[3]))
for key, idx in (('samples', 1), ('instructions', 0), ('local-mem instructions', 3)):
    print('--- by', key)
    for ln, v in sorted(agg.items(), key=lambda x: -x[1][idx])[:top]:
        print(f"{ln[0]}:{ln[1]:<5d} inst {v[0]/max(tot[0],1):6.3f} samples {v[1]/max(tot[1],1):6.3f} localinst {v[3]/max(tot[3],1):6.3f}")
