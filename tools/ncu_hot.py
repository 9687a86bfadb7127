"""Top stall-sample SASS lines of one kernel in an .ncu-rep (source page, read on the CPU box)."""
import csv, io, subprocess, sys
rep, kidx = sys.argv[1], int(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(io.StringIO(out)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "hdr": None, "rows": []}; blocks.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = row
    elif cur is not None and row:
        cur["rows"].append(row)
b = blocks[kidx]
h = b["hdr"]
ia, isrc, iall, ini, iex = h.index("Address"), h.index("Source"), h.index("Warp Stall Sampling (All Samples)"), h.index("Warp Stall Sampling (Not-issued Samples)"), h.index("Instructions Executed")
stall_cols = [i for i, n in enumerate(h) if n.startswith("stall_")]
rows = b["rows"]
tot = sum(int(r[iall]) for r in rows)
print(b["name"][:100], "total samples", tot)
order = sorted(range(len(rows)), key=lambda i: -int(rows[i][iall]))[:top]
for i in sorted(order):
    r = rows[i]
    st = sorted(((int(r[c]), h[c]) for c in stall_cols if r[c] not in ("", "0")), reverse=True)[:3]
    print(f"{i:5d} {int(r[iall]):6d} {100*int(r[iall])/tot:5.1f}%  ex={r[iex]:>8s}  {r[isrc].strip()[:70]:70s} {' '.join(f'{n[6:]}={v}' for v, n in st)}")
