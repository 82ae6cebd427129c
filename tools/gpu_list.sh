#!/bin/bash
# launch list (gpu__time_duration per launch) of one bench_configs case: tools/gpu_list.sh c5
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c ${NLAUNCH:-200} --csv --log-file gpurun_out/list_$1.csv python tools/bench_configs.py $1 > gpurun_out/list_$1.log 2>&1
python - "$1" <<'PY'
import csv, sys, collections
rows = [r for r in csv.reader(open("gpurun_out/list_%s.csv" % sys.argv[1])) if len(r) > 5]
h = rows[0]; ik = h.index("Kernel Name"); iv = h.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[1:]:
    try: v = float(r[iv].replace(",", ""))
    except ValueError: continue
    agg.setdefault(r[ik][:90], []).append(v)
for k, v in agg.items():
    print("%-90s n=%4d mean=%10.1f us total=%10.1f us" % (k, len(v), sum(v) / len(v) / 1e3, sum(v) / 1e3))
PY
