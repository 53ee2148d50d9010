"""Top stall-sample SASS instructions of the first kernel in an .ncu-rep:
python tools/ncu_hot.py rep [N] [kernel-index]"""
import csv, subprocess, sys
rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
# split per kernel
kernels, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "rows": []}
        kernels.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None:
        cur["rows"].append(r)
k = kernels[int(sys.argv[3]) if len(sys.argv) > 3 else 0]
h = k["hdr"]
isamp = h.index("# Samples"); isrc = h.index("Source"); iex = h.index("Instructions Executed")
tot = sum(int(r[isamp] or 0) for r in k["rows"])
totex = sum(int(r[iex] or 0) for r in k["rows"])
print(k["name"], "samples", tot, "warp-inst", totex)
idx = sorted(range(len(k["rows"])), key=lambda i: -int(k["rows"][i][isamp] or 0))[:topn]
for i in sorted(idx):
    r = k["rows"][i]
    print(f"{i:5d} {int(r[isamp]):7d} {100*int(r[isamp])/tot:5.1f}%  ex={r[iex]:>9s}  {r[isrc].strip()}")
