#!/usr/bin/env python
"""Per-launch counters of one bench step, from a light ncu pass, as JSON for profiles/ (read by bench.py).

Capture (one GPU, after the same command exited 0 without ncu):
  ncu --clock-control none --csv --log-file gpurun_out/step_<wl>.csv \
      --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum \
      python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-extras --workload <wl>
Then:  python scripts/ncu_step_metrics.py <workload> gpurun_out/step_<wl>.csv [commit] >> merged into profiles/r02_step_metrics.json

Only the launches of the FIRST mcpt_render of the capture are kept (--steps 1 --warmup 0 --no-extras runs exactly one
before the end-to-end arm): they are the launches between the first k_raygen and the last k_accumulate that follows
it without another k_raygen of pass 0 ... in practice: everything up to the first k_resolve.
"""
import csv
import json
import re
import sys
from collections import defaultdict
from pathlib import Path


def main():
    wl, path = sys.argv[1], sys.argv[2]
    commit = sys.argv[3] if len(sys.argv) > 3 else None
    rows = [r for r in csv.reader(open(path, errors="ignore")) if len(r) > 10]
    hdr = rows[0]
    iid, ik, im, iv, iu = hdr.index("ID"), hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    launches = {}
    for r in rows[1:]:
        i = int(r[iid])
        d = launches.setdefault(i, {"kernel": re.sub(r"^void ", "", r[ik]).replace("mcpt::", "")})
        v = float(r[iv].replace(",", ""))
        unit = r[iu]
        scale = {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "second": 1e3, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9,
                 "byte": 1.0}.get(unit, 1.0)
        d[r[im]] = v * scale
    order = sorted(launches)
    kept = []
    for i in order:
        if launches[i]["kernel"].startswith("k_resolve"):
            break
        kept.append(launches[i])
    out = {"workload": wl, "commit": commit, "source": Path(path).name, "n_launches": len(kept), "launches": []}
    agg = defaultdict(lambda: defaultdict(float))
    n_trav = 0
    for d in kept:
        name = re.sub(r"\(.*", "", d["kernel"])
        stage = ("traverse" if name.startswith("k_traverse") else name)
        rec = {"kernel": name, "ms": d.get("gpu__time_duration.sum", 0.0), "warp_inst": d.get("smsp__inst_executed.sum", 0.0),
               "thread_inst": d.get("smsp__thread_inst_executed.sum", 0.0),
               "issue_active_pct": d.get("smsp__issue_active.avg.pct_of_peak_sustained_active", 0.0),
               "dram_bytes": d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0), "lts_bytes": d.get("lts__t_bytes.sum", 0.0)}
        if stage == "traverse":
            rec["traverse_index"] = n_trav
            n_trav += 1
        out["launches"].append(rec)
        a = agg[stage]
        for k in ("ms", "warp_inst", "thread_inst", "dram_bytes", "lts_bytes"):
            a[k] += rec[k]
        a["issue_ms"] += rec["issue_active_pct"] * rec["ms"]
        a["launches"] += 1
    out["by_stage"] = {}
    for st, a in agg.items():
        out["by_stage"][st] = {"launches": int(a["launches"]), "ms_under_ncu": a["ms"], "warp_inst": a["warp_inst"], "thread_inst": a["thread_inst"],
                               "active_lanes": a["thread_inst"] / a["warp_inst"] if a["warp_inst"] else None,
                               "issue_active_pct_time_weighted": a["issue_ms"] / a["ms"] if a["ms"] else None,
                               "dram_bytes": a["dram_bytes"], "lts_bytes": a["lts_bytes"]}
    trav = [r for r in out["launches"] if "traverse_index" in r]
    if trav:
        first = trav[0]
        out["traverse_first_bounce"] = {"kernel": first["kernel"], "ms_under_ncu": first["ms"], "active_lanes": first["thread_inst"] / max(first["warp_inst"], 1),
                                        "issue_active_pct": first["issue_active_pct"], "dram_bytes": first["dram_bytes"]}
        rest = trav[1:]
        wi = sum(r["warp_inst"] for r in rest)
        if wi:
            out["traverse_later_launches"] = {"launches": len(rest), "active_lanes": sum(r["thread_inst"] for r in rest) / wi,
                                              "issue_active_pct_time_weighted": sum(r["issue_active_pct"] * r["ms"] for r in rest) / max(sum(r["ms"] for r in rest), 1e-9)}
    out.pop("launches")   # the per-launch list stays in the CSV; the JSON holds the aggregates bench.py reads
    dst = Path(__file__).resolve().parent.parent / "profiles" / "r02_step_metrics.json"
    allw = json.loads(dst.read_text()) if dst.exists() else {}
    allw[wl] = out
    dst.write_text(json.dumps(allw, indent=1, sort_keys=True) + "\n")
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
