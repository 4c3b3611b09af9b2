#!/usr/bin/env python
"""profiles/ncu_traffic.json from the ncu summaries of a capture round: dram__bytes_read + write of the two solve kernels per
workload (scripts/gpu_profiles2.sh writes <tag>_ncu_<workload>_{feasible,shifted}_kernel.txt).  usage: ncu_traffic.py <dir> <tag>"""
import json, os, re, sys
d, tag = sys.argv[1], sys.argv[2]
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
out = {}
for w in ("boxer_radial", "pointRobot_sweep", "panda", "boxer_supermarket"):
    det = {}
    for kern in ("feasible", "shifted"):
        p = os.path.join(d, f"{tag}_ncu_{w}_{kern}_kernel.txt")
        if not os.path.isfile(p):
            continue
        for l in open(p):
            m = re.match(r"dram__bytes_(read|write)\.sum\s+(\w+)\s+([0-9.]+)", l)
            if m:
                det[f"{kern}_{m.group(1)}"] = float(m.group(3)) * UNIT[m.group(2)]
    if det:
        out[w] = {"bytes": sum(det.values()), "detail": det,
                  "source": f"profiles/{tag}_ncu_{w}_*_kernel.txt (ncu --set full, each kernel captured alone, default batch)"}
json.dump(out, open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_traffic.json"), "w"), indent=1)
print(json.dumps({k: v["bytes"] for k, v in out.items()}))
