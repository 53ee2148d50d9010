"""Executed warp-instruction histogram by SASS opcode: python tools/ncu_opmix.py rep [kernel-index]"""
import csv, subprocess, sys, collections
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
kernels, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "rows": []}; kernels.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None:
        cur["rows"].append(r)
k = kernels[int(sys.argv[2]) if len(sys.argv) > 2 else 0]
h = k["hdr"]; isrc = h.index("Source"); iex = h.index("Instructions Executed")
hist = collections.Counter()
for r in k["rows"]:
    src = r[isrc].strip()
    toks = src.split()
    if not toks: continue
    op = toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]
    op = op.split(".")[0]
    hist[op] += int(r[iex] or 0)
tot = sum(hist.values())
print(k["name"], "total warp-inst", tot)
for op, c in hist.most_common(30):
    print(f"{op:12s} {c:12d} {100*c/tot:5.1f}%")
