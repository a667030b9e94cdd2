"""Summarises `ncu -i X.ncu-rep --page source --csv --print-source sass` (stall reasons, hottest SASS lines)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
blocks = []; cur = None
for r in rows:
    if r and r[0] == "Kernel Name": cur = {"name": r[1], "hdr": None, "data": []}; blocks.append(cur); continue
    if cur is None: continue
    if r and r[0] == "Address": cur["hdr"] = r; continue
    if cur["hdr"] and len(r) >= len(cur["hdr"]) - 2: cur["data"].append(r)
b = blocks[0]; hdr = b["hdr"]; H = {h: i for i, h in enumerate(hdr)}; data = b["data"]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
g = lambda r, k: int(float(r[H[k]] or 0)) if H[k] < len(r) and r[H[k]] not in ("", "-") else 0
ns = sum(g(r, "# Samples") for r in data); ni = sum(g(r, "Instructions Executed") for r in data)
print(b["name"][:100]); print("samples", ns, "warp-instructions", ni)
for s, v in sorted(((s, sum(g(r, s) for r in data)) for s in stalls), key=lambda kv: -kv[1])[:8]: print(f"  {s:28s} {v:8d} {v/ns:.3f}")
wf = sum(g(r, "L1 Wavefronts Shared") for r in data); wfi = sum(g(r, "L1 Wavefronts Shared Ideal") for r in data)
print("shared wavefronts", wf, "ideal", wfi)
for r in sorted(data, key=lambda r: -g(r, "# Samples"))[:topn]:
    print(f'{g(r,"# Samples"):6d} {g(r,"Instructions Executed"):9d} {r[H["Source"]][:70]:70s} wf {g(r,"L1 Wavefronts Shared"):9d}/{g(r,"L1 Wavefronts Shared Ideal"):9d} ssb {g(r,"stall_short_sb"):5d} mio {g(r,"stall_mio"):5d} bar {g(r,"stall_barrier"):5d} wait {g(r,"stall_wait"):5d} lsb {g(r,"stall_long_sb"):5d} nosel {g(r,"stall_not_selected"):5d}')
