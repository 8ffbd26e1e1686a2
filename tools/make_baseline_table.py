"""Render the measured comparison table of BASELINE.md from profiles/bench_*.json."""
import glob
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load(pattern):
    out = {}
    for p in glob.glob(os.path.join(ROOT, "profiles", pattern)):
        m = re.search(r"_n(\d+)", p)
        try:
            out[int(m.group(1))] = json.loads(open(p).read().strip().splitlines()[-1])
        except Exception:
            pass
    return out


def rows(d):
    r = {"mnist_mlp_sync": {"value": d["value"], "ms": d["ms_per_step"], "e2e": (d.get("e2e") or {}).get("value")}}
    for k, v in d.get("extra", {}).items():
        if "value" in v:
            r[k] = {"value": v["value"], "ms": v["ms_per_step"], "e2e": (v.get("e2e") or {}).get("value"),
                    "tflops": v.get("tflops_per_gpu"), "frac": v.get("fraction_of_measured_bf16_sustained")}
    return r


def main(tag="r1"):
    base = load("bench_baseline_n*.json")
    mine = load(f"bench_b200ps_n*_{tag}.json")
    print("| config | GPUs | baseline ex/s (PyTorch+NCCL) | b200ps ex/s | ratio | b200ps ms/step | b200ps e2e ex/s | baseline e2e ex/s |")
    print("|---|---|---|---|---|---|---|---|")
    for cfg in ("mnist_mlp_sync", "mnist_mlp_async", "sparse_lr", "sparse_wide_deep", "wide_mlp"):
        for n in sorted(mine):
            m = rows(mine[n]).get(cfg)
            b = rows(base[n]).get(cfg) if n in base else None
            if not m:
                continue
            bv = b["value"] if b else None
            print(f"| {cfg} | {n} | {bv:,.0f} | {m['value']:,.0f} | {m['value'] / bv:.2f}x | {m['ms']:.4f} | "
                  f"{(m['e2e'] or 0):,.0f} | {((b or {}).get('e2e') or 0):,.0f} |" if bv else
                  f"| {cfg} | {n} | - | {m['value']:,.0f} | - | {m['ms']:.4f} | {(m['e2e'] or 0):,.0f} | - |")


if __name__ == "__main__":
    import sys
    main(*(sys.argv[1:2]))
