#!/usr/bin/env python
"""Copy the bench lines of a GPU run into profiles/ and fill the @PLACEHOLDERS@ of profiles/README.md.

  python scripts/fill_profiles.py gpurun_out/r29 [gpurun_out/r21_c2_n2.json]
"""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
prefix = sys.argv[1]
n2 = sys.argv[2] if len(sys.argv) > 2 else None


def line(path):
    return json.loads(Path(path).read_text().strip().splitlines()[-1])


src = {"c2": f"{prefix}_c2_default.json", "c1": f"{prefix}_c1.json", "c5": f"{prefix}_c5.json", "mesh": f"{prefix}_mesh.json",
       "ref": f"{prefix}_ref_c2.json"}
d = {k: line(v) for k, v in src.items()}
out = ROOT / "profiles"
(out / "r01_bench_c2.json").write_text(json.dumps(d["c2"]) + "\n")
for k in ("c1", "c5", "mesh"):
    (out / f"r01_bench_{k}.json").write_text(json.dumps(d[k]) + "\n")
(out / "r01_bench_reference_arm.json").write_text(json.dumps(d["ref"]) + "\n")
rep = {}
for k, tag in (("c2", "C2"), ("c1", "C1"), ("c5", "C5"), ("mesh", "MESH")):
    rep[f"@{tag}@"] = f"{d[k]['value']:.0f}"
    rep[f"@{tag}E@"] = f"{d[k]['e2e']['value']:.0f}"
    rep[f"@{tag}MS@"] = f"{d[k]['ms_per_step']:.2f}"
    cb = d[k].get("cpu_baseline") or {}
    rep[f"@{tag}CPU@"] = f"{cb.get('value', 0):.1f}" if cb else "–"
rep["@REF@"] = f"{d['ref']['value']:.1f} (`--impl reference`), {d['c2']['cpu_baseline']['value']:.1f} (`cpu_baseline`)"
if n2:
    x = line(n2)
    (out / "r01_bench_c2_n2.json").write_text(json.dumps(x) + "\n")
    rep["@N2@"], rep["@N2E@"], rep["@N2MS@"] = f"{x['value']:.0f}", f"{x['e2e']['value']:.0f}", f"{x['ms_per_step']:.2f}"
st = d["c2"]["roofline"]["stage_ms_per_step"]
tot = sum(st.values())
rows = ["| stage | ms / step | share |", "|---|---|---|"]
for k, v in st.items():
    if v:
        rows.append(f"| {k} | {v:.2f} | {100 * v / tot:.0f} % |")
rows.append(f"| sum of stages | {tot:.2f} | (step wall: {d['c2']['ms_per_step']:.2f} ms) |")
rep["@STAGES@"] = "\n".join(rows)
p = out / "README.md"
s = p.read_text()
for k, v in rep.items():
    s = s.replace(k, v)
p.write_text(s)
print({k: v for k, v in rep.items() if k != "@STAGES@"})
