#!/bin/bash
# profiles of the final tree: launch list of the default step under ncu, full capture of the two-stage transport launch
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --main-only > gpurun_out/r2af_bench_plain.json 2> gpurun_out/r2af_bench_plain.err; echo "plain rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2af_launches.csv \
   python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --main-only > gpurun_out/r2af_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
python tools/prof_target.py 4096 8 > /dev/null 2>&1; echo "prof_target rc=$?"
ncu --set full --clock-control none --import-source on -k regex:transport_pair_march -c 1 -o gpurun_out/r2af_transport_pair -f \
   python tools/prof_target.py 4096 8 > gpurun_out/r2af_ncu_full_pair.log 2>&1; echo "ncu full pair rc=$?"
python tools/ncu_summary.py gpurun_out/r2af_transport_pair.ncu-rep gpurun_out/r2af_transport_pair_ncu_summary.txt; tail -30 gpurun_out/r2af_transport_pair_ncu_summary.txt
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/r2af_launches.csv")) if len(r)>10]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value"); ui=h.index("Metric Unit")
t=collections.defaultdict(lambda:[0,0.0])
for r in rows[1:]:
    v=float(r[vi].replace(",","")); u=r[ui]
    v = v/1e3 if u in ("ns","nsecond") else (v if u in ("us","usecond") else v*1e3)
    k=r[ki].split("(")[0].replace("void <unnamed>::","")
    t[k][0]+=1; t[k][1]+=v
tot=sum(v[1] for v in t.values())
out=["launch list of `bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --main-only` under ncu (first 400 launches; per-launch times are cold-cache and serialised: shares, not absolutes)"]
for k,(n,us) in sorted(t.items(), key=lambda kv:-kv[1][1]): out.append(f"{k:60s} launches {n:4d}  total {us/1e3:9.3f} ms  avg {us/n:9.1f} us  share {us/tot:6.3f}")
open("gpurun_out/r2af_launches_summary.txt","w").write("\n".join(out)+"\n"); print("\n".join(out))
PY
