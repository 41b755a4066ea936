#!/usr/bin/env python
"""Per-source-line summary of an ncu report: ncu -i X.ncu-rep --page source --print-source cuda,sass --csv > f.csv;
   python scripts/ncu_lines.py f.csv FRAMES [min_share]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
F = float(sys.argv[2]); thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.004
def num(x):
    try: return float(x)
    except ValueError: return 0.0
fil = None; hdr = None; cur = None
agg = collections.OrderedDict()
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fil = r[1].split("/")[-1]; continue
    if len(r) > 4 and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr) - 2: continue
    if r[0] != "":
        cur = (fil, int(r[0]), r[1].strip()[:90]); agg.setdefault(cur, collections.Counter()); continue
    if r[2] in ("", "..."): continue
    d = dict(zip(hdr[2:], r[2:])); a = agg[cur]
    a["instr"] += num(d["Instructions Executed"]) / F; a["samples"] += num(d["# Samples"])
    a["wf"] += num(d["L1 Wavefronts Shared"]) / F
    for k in d:
        if k.startswith("stall_") and "Not Issued" not in k: a[k] += num(d[k])
tot = sum(a["samples"] for a in agg.values()); ti = sum(a["instr"] for a in agg.values())
print("total instr/frame %.0f samples %d" % (ti, tot))
for k, a in agg.items():
    if a["samples"] / tot < thr and a["instr"] < 20: continue
    st = sorted(((v, s[6:]) for s, v in a.items() if s.startswith("stall_")), reverse=True)[:3]
    print("%-18s %4d  ins %6.1f  smp %5.2f%%  wf %5.1f  %-44s | %s" % (k[0], k[1], a["instr"], 100 * a["samples"] / tot, a["wf"],
          " ".join("%s:%.0f%%" % (s, 100 * v / max(a["samples"], 1)) for v, s in st), k[2][:70]))
