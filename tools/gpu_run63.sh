timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_c2_sb.csv python tools/prof.py --workload c2 --lin 1 --pcg 20000 --gn 3 > gpurun_out/ncu_sb.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open('gpurun_out/launches_c2_sb.csv')) if len(r)>5]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
agg=collections.defaultdict(list)
for r in rows[1:]:
    try: agg[r[ki][:60]].append(float(r[vi].replace(',','')))
    except: pass
for k,v in sorted(agg.items(), key=lambda kv:-sum(kv[1]))[:10]: print(f"{k:60s} n={len(v):4d} avg={sum(v)/len(v)/1e3:9.1f} us")
PY
echo "--- c2 50 GN adaptive"; timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 20000 --gn 50 2>&1 | grep "optimize" | cut -c1-400
timeout 900 python -m pytest tests -m gpu -x -q -k "resident or full_size or reproducible" 2>&1 | tail -2
