#!/usr/bin/env python
"""figs/perf.svg -- the counterpart of the reference's figs/perf.png (README.md:8): join time against the
shuffle window W, both axes log2, one curve per algorithm.

  python experiments/plot_perf.py [--in experiments/results/w_sweep.jsonl] [--out figs/perf.svg]

GPU curves come from experiments/w_sweep.py (this library on a B200); the CPU curves are the reference's own
published medians (experiments/published_cpu.json, another machine) and are drawn dashed.  No plotting
library is available in this image, so the SVG is written directly.
"""
import argparse
import json
import math
import os

HERE = os.path.dirname(os.path.abspath(__file__))
COLORS = ["#d62728", "#1f77b4", "#2ca02c", "#9467bd", "#8c564b", "#e377c2", "#7f7f7f", "#bcbd22"]
GPU_NAMES = {"nocc": "NoCC global table (B200)", "atomic": "Atomic global table (B200)",
             "radix": "Radix join (B200)", "adaptive": "Adaptive (B200)", "sortmerge": "Sort-merge (B200)"}


def load_gpu(path):
    curves = {}
    for l in open(path):
        l = l.strip()
        if not l.startswith("{"):
            continue
        o = json.loads(l)
        if o.get("dataDistr") != "local_shuffle":
            continue
        curves.setdefault(GPU_NAMES.get(o["algo"], o["algo"]), []).append(
            (math.log2(o["shuffleRange"]), o["hashBuildTimeInMicroseconds"] / 1000.0))
    return {k: sorted(v) for k, v in curves.items()}


def load_cpu(path, family):
    d = json.load(open(path))[family]
    xs = d["log2_W"]
    return {k: [(x, us / 1000.0) for x, us in zip(xs, v)] for k, v in d.items() if not k.startswith("_") and k != "log2_W"}


def load_measured_cpu(path):
    """bench.py --cpu-curves: the reference's CPU code timed on the GPU box's own host cores."""
    curves = {}
    for l in open(path):
        l = l.strip()
        if l.startswith("{"):
            o = json.loads(l)
            curves.setdefault(o["algo"], []).append((math.log2(o["shuffleRange"]), o["hashBuildTimeInMicroseconds"] / 1000.0))
    return {k: sorted(v) for k, v in curves.items()}


def svg(gpu, cpu, title, out, note=""):
    W, H, L, R, T, B = 900, 560, 70, 270, 40, 55
    pts = [p for c in list(gpu.values()) + list(cpu.values()) for p in c]
    x0, x1 = 0.0, max(p[0] for p in pts)
    y0 = math.floor(math.log2(min(p[1] for p in pts)))
    y1 = math.ceil(math.log2(max(p[1] for p in pts)))

    def X(x):
        return L + (x - x0) / (x1 - x0) * (W - L - R)

    def Y(ms):
        return H - B - (math.log2(ms) - y0) / (y1 - y0) * (H - T - B)

    o = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{W}" height="{H}" font-family="sans-serif" font-size="12">',
         f'<rect width="{W}" height="{H}" fill="white"/>',
         f'<text x="{L}" y="22" font-size="15">{title}</text>']
    for gx in range(0, int(x1) + 1, 3):
        o.append(f'<line x1="{X(gx):.1f}" y1="{T}" x2="{X(gx):.1f}" y2="{H - B}" stroke="#e6e6e6"/>')
        o.append(f'<text x="{X(gx):.1f}" y="{H - B + 16}" text-anchor="middle">2^{gx}</text>')
    for gy in range(y0, y1 + 1):
        o.append(f'<line x1="{L}" y1="{Y(2.0 ** gy):.1f}" x2="{W - R}" y2="{Y(2.0 ** gy):.1f}" stroke="#e6e6e6"/>')
        lab = f"{2.0 ** gy:g}"
        o.append(f'<text x="{L - 6}" y="{Y(2.0 ** gy) + 4:.1f}" text-anchor="end">{lab}</text>')
    o.append(f'<rect x="{L}" y="{T}" width="{W - L - R}" height="{H - T - B}" fill="none" stroke="black"/>')
    o.append(f'<text x="{(L + W - R) / 2}" y="{H - 12}" text-anchor="middle">shuffle window W (tuples)</text>')
    o.append(f'<text transform="translate(16,{(T + H - B) / 2}) rotate(-90)" text-anchor="middle">'
             f'build + probe time (ms, log2), |R| = |S| = 2^27</text>')
    ly = T + 10
    for i, (name, c, dash) in enumerate([(k, v, False) for k, v in gpu.items()] + [(k, v, True) for k, v in cpu.items()]):
        col = COLORS[i % len(COLORS)]
        path = " ".join(f"{'M' if j == 0 else 'L'}{X(x):.1f},{Y(y):.1f}" for j, (x, y) in enumerate(c))
        da = ' stroke-dasharray="6,4"' if dash else ""
        o.append(f'<path d="{path}" fill="none" stroke="{col}" stroke-width="2"{da}/>')
        for x, y in c:
            o.append(f'<circle cx="{X(x):.1f}" cy="{Y(y):.1f}" r="2.5" fill="{col}"/>')
        o.append(f'<line x1="{W - R + 12}" y1="{ly}" x2="{W - R + 42}" y2="{ly}" stroke="{col}" stroke-width="2"{da}/>')
        o.append(f'<text x="{W - R + 48}" y="{ly + 4}">{name}</text>')
        ly += 20
    o.append(f'<text x="{W - R + 12}" y="{ly + 8}" font-size="10">dashed: CPU (reference code): published logs (2016),</text>')
    o.append(f'<text x="{W - R + 12}" y="{ly + 20}" font-size="10">"this host" = timed on the GPU box {note}</text>')
    o.append("</svg>")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    open(out, "w").write("\n".join(o) + "\n")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--in", dest="inp", default=os.path.join(HERE, "results", "w_sweep.jsonl"))
    ap.add_argument("--out", default=os.path.join(HERE, "..", "figs", "perf.svg"))
    ap.add_argument("--family", default="build_probe", choices=["build_probe", "build_only"])
    ap.add_argument("--cpu-measured", default="", help="bench.py --cpu-curves output: the reference's CPU code on this box")
    a = ap.parse_args()
    gpu = load_gpu(a.inp)
    cpu = load_cpu(os.path.join(HERE, "published_cpu.json"), a.family)
    note = ""
    if a.cpu_measured and os.path.exists(a.cpu_measured):
        meas = load_measured_cpu(a.cpu_measured)
        cpu.update(meas)
        first = json.loads(open(a.cpu_measured).readline())
        note = f"(|R| = {first.get('rSize')}, {first.get('cores')} cores)"
    svg(gpu, cpu, "Hash join vs data locality: B200 library and the reference's CPU runs", a.out, note)
    print("wrote", os.path.normpath(a.out), {k: len(v) for k, v in gpu.items()})


if __name__ == "__main__":
    main()
