"""DRAM traffic per launch of the dominant kernels, measured (not estimated): runs the smallest driver of every
configuration under `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum` and writes
profiles/traffic.json (what bench.py reports as roofline.traffic) and a per-kernel table.

    python tools/measure_traffic.py            (on the GPU box; every driver runs once WITHOUT ncu first)
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
RUNS = [("k1tc", [sys.executable, "tools/tc_prof_driver.py", "2"], {}),
        ("configs", [sys.executable, "tools/bench_configs.py"], {"PYFFT_BENCH_ONCE": "1"})]
METRICS = "dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum"


def main():
    os.makedirs(OUT, exist_ok=True)
    table = {}
    for tag, cmd, env in RUNS:
        e = dict(os.environ, **env)
        plain = subprocess.run(cmd, cwd=ROOT, env=e, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if plain.returncode != 0:
            print("plain run of %s failed:\n%s" % (tag, plain.stdout[-2000:]))
            return 1
        log = os.path.join(OUT, "traffic_%s.csv" % tag)
        r = subprocess.run(["ncu", "--metrics", METRICS, "--clock-control", "none", "--csv", "--log-file", log] + cmd,
                           cwd=ROOT, env=e, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            print("ncu run of %s failed:\n%s" % (tag, r.stdout[-2000:]))
            return 1
        rows = [x for x in csv.reader(open(log)) if len(x) > 10]
        h = rows[0]
        ik, im, iv, iu = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
        iid = h.index("ID")
        per = {}
        for x in rows[1:]:
            v = float(x[iv].replace(",", ""))
            u = x[iu]
            v *= {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1e3, "us": 1.0, "ns": 1e-3, "s": 1e6,
                  "msecond": 1e3, "usecond": 1.0, "nsecond": 1e-3, "second": 1e6}.get(u, 1.0)
            per.setdefault((x[iid], x[ik].split("(")[0].split("<")[0]), {})[x[im]] = v
        for (_, name), m in per.items():
            t = table.setdefault(tag + ":" + name, dict(launches=0, read=0.0, write=0.0, us=0.0))
            t["launches"] += 1
            t["read"] += m.get("dram__bytes_read.sum", 0.0)
            t["write"] += m.get("dram__bytes_write.sum", 0.0)
            t["us"] += m.get("gpu__time_duration.sum", 0.0)
    lines = ["kernel (driver:name)                                   launches   read MB/launch  write MB/launch   us/launch"]
    for k, t in sorted(table.items(), key=lambda kv: -kv[1]["us"]):
        n = t["launches"]
        lines.append("%-55s %8d %16.2f %16.2f %11.1f" % (k[:55], n, t["read"] / n / 1e6, t["write"] / n / 1e6, t["us"] / n))
    open(os.path.join(OUT, "traffic_table.txt"), "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[:40]))
    tc = table.get("k1tc:welch_tc_kernel") or table.get("k1tc:pyfft::welch_tc_kernel")
    if tc:
        n = tc["launches"]
        out = {"welch_tc_kernel (tcgen05)": {"dram_bytes_read": int(tc["read"] / n), "dram_bytes_write": int(tc["write"] / n),
                                             "source": "tools/measure_traffic.py (ncu --metrics dram__bytes_*.sum, mean of %d launches of "
                                                       "tools/tc_prof_driver.py, 2 x 2^26 fp32 samples)" % n}}
        json.dump(out, open(os.path.join(OUT, "traffic.json"), "w"), indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
