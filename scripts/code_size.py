"""SASS instruction count of solve_kernel per source region (from nvdisasm --print-line-info of a solver .so)."""
import re, sys, subprocess, collections, os, tempfile
so = sys.argv[1]
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
open(sys.argv[2] if len(sys.argv) > 2 else "/tmp/dis.txt", "w").write(dis)
cnt = collections.Counter(); cur = None; infunc = False; total = 0
for l in dis.splitlines():
    if l.startswith("\t.text.") or l.startswith(".text."): infunc = "solve_kernel" in l
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2)))
    if infunc and re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+\S", l):
        total += 1
        if cur: cnt[(cur[0], cur[1] // 20 * 20)] += 1
print("total SASS instructions in solve_kernel:", total, "=", total * 16 // 1024, "KB")
for (f, ln), c in sorted(cnt.items(), key=lambda x: -x[1])[:30]:
    print(f"{f}:{ln}-{ln+19}  {c}  {c/total:.3f}")
