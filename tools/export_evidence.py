#!/usr/bin/env python
"""Turns what tools/gpu_evidence_1gpu.sh left in gpurun_out/ (ev_*) into the tracked summaries under profiles/ (r02_*)."""
import csv
import json
import re
import shutil
import subprocess
import sys

G, P = "gpurun_out/ev_", "profiles/r02_"


def launches():
    rows = list(csv.reader(open(G + "launches_traffic.csv", errors="replace")))
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    h = rows[start]
    kn, mn, mv, idc, gs = (h.index(x) for x in ("Kernel Name", "Metric Name", "Metric Value", "ID", "Grid Size"))
    d, order = {}, []
    for r in rows[start + 1:]:
        if len(r) <= mv:
            continue
        if r[idc] not in d:
            d[r[idc]] = {"k": r[kn], "grid": r[gs]}
            order.append(r[idc])
        d[r[idc]][r[mn]] = float(r[mv].replace(",", ""))

    def short(n):
        return re.sub(r"\(.*", "", n).replace("void ", "").replace("<unnamed>::", "").replace("unnamed>::", "").replace("ftf::", "")

    step = order[:len(order) // 2]          # the capture holds two identical steps
    with open(P + "launches_inception_b64.csv", "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["launch", "kernel", "grid", "time_us", "dram_read_MB", "dram_write_MB"])
        for k in step:
            v = d[k]
            w.writerow([k, short(v["k"]), v["grid"], "%.2f" % (v["gpu__time_duration.sum"] / 1e3), "%.2f" % (v["dram__bytes_read.sum"] / 1e6),
                        "%.2f" % (v["dram__bytes_write.sum"] / 1e6)])
    tr = sum(d[k]["dram__bytes_read.sum"] for k in step)
    tw = sum(d[k]["dram__bytes_write.sum"] for k in step)
    tt = sum(d[k]["gpu__time_duration.sum"] for k in step)
    conv = [k for k in step if "conv_tc_kernel" in d[k]["k"] or "conv_halo" in d[k]["k"]]
    ct = sum(d[k]["gpu__time_duration.sum"] for k in conv)
    t = json.load(open("profiles/traffic.json"))
    t["inception"].update({"kernel": "all %d launches of one batch-64 forward pass (%d tensor-core convolutions = %.2f of the kernel time)" % (len(step), len(conv), ct / tt),
                           "dram_bytes_per_launch": int(tr + tw), "dram_read_bytes": int(tr), "dram_write_bytes": int(tw)})
    json.dump(t, open("profiles/traffic.json", "w"), indent=1)
    print("launch list: %d launches, convolutions %.3f of %.1f us, DRAM %.1f MB read %.1f MB written" % (len(step), ct / tt, tt / 1e3, tr / 1e6, tw / 1e6))


def main():
    launches()
    for x in "ABC":
        subprocess.run([sys.executable, "tools/ncu_summary.py", G + "conv_full_%s.ncu-rep" % x, P + "ncu_conv_full_%s.csv" % x], check=True)
    subprocess.run([sys.executable, "tools/ncu_summary.py", G + "stem_pools_full.ncu-rep", P + "ncu_stem_pools_full.csv"], check=True)
    for n in ("inception", "hpt", "normalize", "jpeg", "reference"):
        line = open(G + "bench_%s.json" % n).read().strip().splitlines()[-1]
        json.loads(line)
        open(P + "bench_%s.json" % n, "w").write(line + "\n")
    out, on = [], False
    for l in open(G + "timeline.txt"):
        if l.startswith("timeline batch"):
            if out:
                break
            on = True
        if on:
            out.append(l)
            if l.startswith("sum of"):
                break
    open(P + "timeline_final.txt", "w").writelines(out)
    shutil.copy(G + "plan.txt", P + "conv_plan.txt")


if __name__ == "__main__":
    main()
