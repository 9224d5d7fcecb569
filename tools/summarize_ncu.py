"""Turns the two ncu CSV passes over tools/profile_step.py into the committed evidence under profiles/:

  python tools/summarize_ncu.py <tag> <launches.csv> <traffic.csv> <profile_step_classes.json> <plain_run_ms> <bench.json>

  launches.csv : ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv
  traffic.csv  : ncu --profile-from-start off --metrics dram__bytes_read.sum,dram__bytes_write.sum --csv
  classes.json : written by tools/profile_step.py (kernel class of every launch, in launch order)

Writes profiles/<tag>_ncu_launches.csv (copy), profiles/<tag>_ncu_launches_summary.md, profiles/<tag>_traffic.json and
profiles/traffic.json (dram bytes per launch per kernel class -- what bench.py reports as roofline.traffic)."""
import collections
import csv
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def rows_of(path):
    allr = list(csv.reader(open(path)))
    hdr = [r for r in allr if len(r) > 5 and r[0] == "ID"][0]
    return hdr, [r for r in allr if len(r) > 5 and r[0].isdigit()]


def main():
    tag, launches_csv, traffic_csv, classes_json, plain_ms, bench_json = sys.argv[1:7]
    meta = json.load(open(classes_json))
    seq = meta["sequence"]  # class of every launch; the timings in this file were taken under ncu and are not used
    bench = json.load(open(bench_json))
    cls_ms = {c: {"ms": v["ms_per_step"]} for c, v in bench["kernel_classes"].items()}
    hdr, rows = rows_of(launches_csv)
    ik, ig, iv = hdr.index("Kernel Name"), hdr.index("Grid Size"), hdr.index("Metric Value")
    by_kernel, by_class = collections.OrderedDict(), collections.Counter()
    total = 0.0
    for i, r in enumerate(rows):
        us = float(r[iv].replace(",", "")) / 1000.0
        name = r[ik].split("(")[0].replace("void ", "").replace("dmx::", "")
        a = by_kernel.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += us
        by_class[seq[i] if i < len(seq) else "?"] += us
        total += us
    out = []
    out.append("# ncu launch list, %s: one batch of 16 segments (htdemucs 4s, bf16 mode)\n" % tag)
    out.append("Command (B200, `gpurun`): `ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none "
               "--csv --log-file X python tools/profile_step.py` (the plain run of the same command exited 0 first: %s ms "
               "device time for the 16 segments).  Per-launch times under ncu are serialised and cold-cache, and the second "
               "stream of the time branch cannot overlap: compare SHARES.\n" % plain_ms)
    out.append("%d kernel launches, %.1f ms summed.\n" % (len(rows), total / 1000.0))
    out.append("| kernel | launches | total us | share |\n|---|---:|---:|---:|")
    for name, (n, us) in sorted(by_kernel.items(), key=lambda kv: -kv[1][1]):
        out.append("| `%s` | %d | %.1f | %.1f %% |" % (name, n, us, 100.0 * us / total))
    out.append("\nBy kernel class (the classes of bench.py's `kernel_classes`; class of each launch from "
               "`dmx_profile_sequence`):\n")
    out.append("| class | ncu us (batch of 16) | ncu share | bench.py live CUDA-event ms / step (31 segments) | live share |\n|---|---:|---:|---:|---:|")
    live_total = sum(v["ms"] for v in cls_ms.values())
    for c, us in by_class.most_common():
        lm = cls_ms.get(c, {}).get("ms", 0.0)
        out.append("| %s | %.1f | %.1f %% | %.2f | %.1f %% |" % (c, us, 100.0 * us / total, lm, 100.0 * lm / max(live_total, 1e-9)))
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    shutil.copy(launches_csv, os.path.join(ROOT, "profiles", "%s_ncu_launches.csv" % tag))
    open(os.path.join(ROOT, "profiles", "%s_ncu_launches_summary.md" % tag), "w").write("\n".join(out) + "\n")

    # ---- dram traffic per class ----
    hdr, rows = rows_of(traffic_csv)
    iid, im, iv = hdr.index("ID"), hdr.index("Metric Name"), hdr.index("Metric Value")
    iu = hdr.index("Metric Unit")
    per_launch = collections.defaultdict(float)
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for r in rows:
        if r[im].startswith("dram__bytes_"):
            per_launch[int(r[iid])] += float(r[iv].replace(",", "")) * scale.get(r[iu], 1.0)
    ids = sorted(per_launch)
    tr = collections.OrderedDict()
    for j, lid in enumerate(ids):
        c = seq[j] if j < len(seq) else "?"
        e = tr.setdefault(c, {"launches_per_batch": 0, "dram_bytes_per_batch": 0.0})
        e["launches_per_batch"] += 1
        e["dram_bytes_per_batch"] += per_launch[lid]
    for c, e in tr.items():
        e["dram_bytes_per_launch"] = e["dram_bytes_per_batch"] / e["launches_per_batch"]
    json.dump(tr, open(os.path.join(ROOT, "profiles", "%s_traffic.json" % tag), "w"), indent=1)
    tj = {c: e["dram_bytes_per_launch"] for c, e in tr.items()}
    tj["_segments_per_launch"] = 16  # tools/profile_step.py profiles one batch of 16 segments, what bench.py launches
    json.dump(tj, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
    shutil.copy(traffic_csv, os.path.join(ROOT, "profiles", "%s_ncu_traffic.csv" % tag))
    print("\n".join(out[-14:]))
    print(json.dumps({c: round(e["dram_bytes_per_launch"] / 1e6, 1) for c, e in tr.items()}))


if __name__ == "__main__":
    main()
