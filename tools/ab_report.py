#!/usr/bin/env python
"""tools/ab_report.py PREFIX VARIANT...: table of the A/B outputs written by tools/ab.sh."""
import json, re, sys
pre, vs = sys.argv[1], sys.argv[2:]
def load(v, suf=""):
    out = {}
    try:
        lines = [json.loads(l) for l in open(f"gpurun_out/{pre}_{v}{suf}.json") if l.startswith("{")]
    except FileNotFoundError:
        return None
    for l in lines:
        if l.get("kernel_only"): out["ko_ms"] = l["ms_per_step"]
        else: out.update(value=l["value"], ms=l["ms_per_step"], e2e=l["e2e"]["value"], frac=l["roofline"]["frac"])
    for l in open(f"gpurun_out/{pre}_{v}{suf}.err"):
        if l.startswith("per-kernel"): out["pk"] = json.loads(l.split(":", 1)[1])
        if l.startswith("e2e ms"): out["e2e_each"] = l.strip()
    return out
for suf in ("", "_4k"):
    print("==", "1080p8" if not suf else "4k10qm")
    for v in vs:
        d = load(v, suf)
        if not d: continue
        pk = d.get("pk", {})
        print(f"{v:10s} value {d.get('value',0):9.0f} ms {d.get('ms',0):.4f} ko_ms {d.get('ko_ms',0):.4f} frac {d.get('frac',0):.3f} e2e {d.get('e2e',0):7.0f} | " +
              " ".join(f"{k}:{x['us']:.1f}" for k, x in pk.items()))
        if "e2e_each" in d: print("           ", d["e2e_each"])
