"""Static SASS opcode mix of solve_kernel per source-line range of rmpc_ipm.cuh.  usage: phase_mix.py <dis.txt> name:lo-hi ..."""
import re, sys, collections
rng = []
for a in sys.argv[2:]:
    n, r = a.split(':'); lo, hi = r.split('-'); rng.append((n, int(lo), int(hi)))
cur = None; infunc = False; ops = collections.defaultdict(collections.Counter)
for l in open(sys.argv[1]):
    if '.text.' in l: infunc = 'solve_kernel' in l
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2)))
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)', l)
    if m and infunc and cur and cur[0] == 'rmpc_ipm.cuh':
        for n, lo, hi in rng:
            if lo <= cur[1] <= hi: ops[n][m.group(3).split('.')[0]] += 1
for n, c in ops.items():
    print(n, sum(c.values()), c.most_common(14))
