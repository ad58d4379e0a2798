# round 2, run K: per-launch durations of the env loop (balance on/off)
mkdir -p gpurun_out/r2k
ncu --metrics gpu__time_duration.sum --clock-control none -s 700 -c 40 --csv --log-file gpurun_out/r2k/launches_bal.csv python tools/phase_profile.py 160 20 > gpurun_out/r2k/ncu_bal.log 2>&1
GGL_BALANCE=0 ncu --metrics gpu__time_duration.sum --clock-control none -s 520 -c 40 --csv --log-file gpurun_out/r2k/launches_nobal.csv python tools/phase_profile.py 160 20 > gpurun_out/r2k/ncu_nobal.log 2>&1
python - <<'PY'
import csv, collections
for f in ("bal","nobal"):
    rows=list(csv.reader(open("gpurun_out/r2k/launches_%s.csv"%f)))
    hdr=[i for i,r in enumerate(rows) if r and r[0]=="ID"][0]
    h=rows[hdr]; kn=h.index("Kernel Name"); mv=h.index("Metric Value"); mu=h.index("Metric Unit")
    agg=collections.defaultdict(list)
    for r in rows[hdr+1:]:
        if len(r)>mv:
            v=float(r[mv].replace(",","")); u=r[mu]
            v = v/1000 if u=="ns" else (v*1000 if u=="ms" else v)
            agg[r[kn].split("(")[0]].append(v)
    print(f, {k:(len(v), round(sum(v)/len(v),1)) for k,v in agg.items()})
PY
