"""Rank SASS instructions of an ncu source-page CSV by warp-stall samples.
usage: ncu -i X.ncu-rep --page source --csv > x.csv ; python scripts/ncu_hot.py x.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 25
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "data": []}
        blocks.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None and len(r) == len(cur["hdr"]):
        cur["data"].append(r)
b = blocks[0]
hdr, data = b["hdr"], b["data"]
idx = {h: i for i, h in enumerate(hdr)}
def iv(r, h):
    try: return int(r[idx[h]])
    except Exception: return 0
tot = sum(iv(r, "# Samples") for r in data)
print(b["name"], "| launches in file:", len(blocks), "| samples:", tot, "| instructions:", len(data))
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {h: sum(iv(r, h) for r in data) for h in stalls}
print("stall mix:", ", ".join(f"{k[6:]} {v / max(tot, 1):.2f}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:7]))
for r in sorted(data, key=lambda r: -iv(r, "# Samples"))[:N]:
    st = {h: iv(r, h) for h in stalls}
    main = max(st, key=st.get)
    print(f"{iv(r, '# Samples'):7d} {iv(r, '# Samples') / max(tot, 1):6.3f} {main[6:]:14s} {r[idx['Source']][:100]}")
