"""profiles/traffic.json from an ncu --set full raw-page summary (scripts/gpu_profile.sh writes <tag>_full_summary.csv):
dram__bytes_read.sum + dram__bytes_write.sum per launch, largest launch of every kernel (resize: the level-1 launch).
usage: python scripts/make_traffic.py profiles/r01_ncu_full_v6.csv "<note>" > profiles/traffic.json"""
import csv, json, sys

rows = list(csv.reader(open(sys.argv[1])))
h, units = rows[0], rows[1]
ik, ir, iw, ig = h.index("Kernel Name"), h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum"), h.index("Grid Size")
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
out = {}
for r in rows[2:]:
    if len(r) <= iw:
        continue
    name = r[ik].split("(")[0].split("<")[0].split("::")[-1].split()[-1]
    b = float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]]
    out[name] = max(out.get(name, 0), int(b))
out["_note"] = sys.argv[2] if len(sys.argv) > 2 else "dram bytes per launch from " + sys.argv[1]
print(json.dumps(out, indent=1))
