"""Summarises an ncu launch list (profiles/ncu_launches.sh): per kernel the mean duration and DRAM bytes, launches per
step, share of the step; writes the traffic record bench.py reads (profiles/r02_traffic.json).
    python profiles/ncu_extract.py gpurun_out/<tag>_launches.csv <batch> [--write]"""
import collections
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def load(path):
    rows = list(csv.reader(open(path, errors="replace")))
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr = rows[start]
    ki, mi, ni, gi, ii = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name"), hdr.index("Grid Size"), hdr.index("ID")
    launches = collections.OrderedDict()
    for r in rows[start + 1:]:
        if len(r) <= mi:
            continue
        try:
            v = float(r[mi].replace(",", ""))
        except ValueError:
            continue
        d = launches.setdefault(r[ii], {"name": r[ki].split("(")[0].replace("void ", ""), "grid": r[gi]})
        d[r[ni]] = v
    return list(launches.values())


def main():
    path, batch = sys.argv[1], int(sys.argv[2])
    L = [l for l in load(path) if "at::" not in l["name"] and "nccl" not in l["name"].lower()]
    steps = sum(1 for l in L if l["name"].startswith(("k_path_v4", "k_path_v3", "k_path_v2", "k_path_v<")))
    agg = collections.OrderedDict()
    for l in L:
        a = agg.setdefault(l["name"], {"n": 0, "ns": 0.0, "rd": 0.0, "wr": 0.0, "grid": l["grid"]})
        a["n"] += 1; a["ns"] += l.get("gpu__time_duration.sum", 0.0); a["rd"] += l.get("dram__bytes_read.sum", 0.0); a["wr"] += l.get("dram__bytes_write.sum", 0.0)
    tot_ns = sum(a["ns"] for a in agg.values())
    tot_b = sum(a["rd"] + a["wr"] for a in agg.values())
    print("%d steps, %d launches (%.1f per step), %.1f us per frame under ncu, DRAM %.1f MB per frame" % (
        steps, len(L), len(L) / steps, tot_ns / steps / batch / 1e3, tot_b / steps / batch / 1e6))
    print("%-52s %5s %10s %8s %10s %10s  %s" % ("kernel", "n/st", "us/launch", "share", "MB rd/fr", "MB wr/fr", "grid"))
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ns"]):
        print("%-52s %5.1f %10.1f %7.1f%% %10.2f %10.2f  %s" % (k[:52], a["n"] / steps, a["ns"] / a["n"] / 1e3, 100 * a["ns"] / tot_ns,
                                                               a["rd"] / steps / batch / 1e6, a["wr"] / steps / batch / 1e6, a["grid"]))
    if "--write" in sys.argv:
        import bench
        rec = {"src_sha": bench.source_sha(), "batch": batch, "source": os.path.basename(path), "steps_in_capture": steps,
               "dram_bytes_per_frame": tot_b / steps / batch, "launches_per_step": len(L) / steps, "kernels": {}}
        for k, a in agg.items():
            key = k.split("<")[0]
            e = rec["kernels"].setdefault(key, {"dram_read": 0.0, "dram_write": 0.0, "ns": 0.0, "launches_per_step": 0.0, "grid": a["grid"]})
            # per step (all launches of that kernel family in one step)
            e["dram_read"] += a["rd"] / steps; e["dram_write"] += a["wr"] / steps; e["ns"] += a["ns"] / steps; e["launches_per_step"] += a["n"] / steps
        with open(os.path.join(ROOT, "profiles", "r02_traffic.json"), "w") as f:
            json.dump(rec, f, indent=1)
        print("wrote profiles/r02_traffic.json for sources", rec["src_sha"])


if __name__ == "__main__":
    main()
