#!/usr/bin/env python
"""SASS listing of the Riccati sweep (rmpc_ipm.cuh: sweep()) of the hot kernel solve_kernel<1> of one solver library, from
nvdisasm --print-line-info, with the opcode mix.  usage: python scripts/sass_sweep.py <lib.so>   (run here, no GPU)"""
import collections, os, re, subprocess, sys, tempfile
so = os.path.abspath(sys.argv[1])
src = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "robot_mpcs_b200", "csrc", "rmpc_ipm.cuh")).read().splitlines()
lo = next(i for i, l in enumerate(src) if "RMPC_HD void sweep(" in l) + 1
hi = next(i for i, l in enumerate(src) if "forward roll-out" in l and i > lo)
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", so], cwd=tmp, stdout=subprocess.DEVNULL)
cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cub)], capture_output=True, text=True).stdout.splitlines()
fn = cur = None
rows = []
for l in dis:
    m = re.match(r"\s*\.text\.(\S+):", l)
    if m: fn = m.group(1)
    if not fn or "solve_kernelILi1" not in fn: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2)))
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m: rows.append((int(m.group(1), 16), cur, m.group(2).strip()))
idx = [i for i, (a, c, t) in enumerate(rows) if c and c[0] == "rmpc_ipm.cuh" and lo <= c[1] <= hi]
a, b = min(idx), max(idx)
body = rows[a:b + 1]
ops = collections.Counter((t.split()[1] if t.startswith("@") else t.split()[0]).split(".")[0] for _, _, t in body)
print(f"# {os.path.basename(so)}: sweep() = rmpc_ipm.cuh:{lo}-{hi}; {len(body)} SASS instructions for one stage (all three phases), of {len(rows)} in solve_kernel<1>")
print("# opcode mix:", ", ".join(f"{k} {v}" for k, v in ops.most_common(16)))
for addr, c, t in body:
    print(f"{addr:06x}  {(c[0][:14] + ':' + str(c[1])) if c else '':20s} {t}")
