"""Turn an `ncu --metrics gpu__time_duration.sum --csv` log into the files kept under profiles/:
a compact per-launch CSV and a markdown table of time per kernel, plus the breakdown of one decode step
(the launches between the last two argmax kernels).

    python tools/summarize_launches.py gpurun_out/bench_launches.csv profiles/r1_bench_launches_window.csv > summary.md
"""

from __future__ import annotations

import collections
import csv
import sys


def read(path):
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    r = csv.reader(lines)
    hdr = next(r)
    ki, bi, gi, ui, vi = (hdr.index(n) for n in ("Kernel Name", "Block Size", "Grid Size", "Metric Unit", "Metric Value"))
    scale = {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}
    return [(x[ki], x[gi], x[bi], float(x[vi].replace(",", "")) * scale.get(x[ui], 1.0)) for x in r]


def table(rows, top=20):
    agg = collections.defaultdict(lambda: [0, 0.0])
    for k, _, _, ns in rows:
        agg[k[:110]][0] += 1
        agg[k[:110]][1] += ns
    total = sum(v[1] for v in agg.values())
    out = ["| launches | ms | share | kernel |", "|---|---|---|---|"]
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        out.append(f"| {n} | {ns / 1e6:.3f} | {100 * ns / total:.1f}% | `{k}` |")
    return total, "\n".join(out)


def main():
    src, dst = sys.argv[1], sys.argv[2]
    rows = read(src)
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["id", "kernel", "grid", "block", "gpu__time_duration.sum_ns"])
        for i, (k, g, b, ns) in enumerate(rows):
            w.writerow([i, k[:160], g, b, int(ns)])
    total, tab = table(rows)
    print(f"Total: {len(rows)} launches, {total / 1e6:.1f} ms.\n")
    print(tab)
    ours = [r for r in rows if "ia::" in r[0][:12]]
    print(f"\nOur kernels in the window: {len(ours)} launches, {sum(r[3] for r in ours) / 1e6:.2f} ms "
          f"({100 * sum(r[3] for r in ours) / total:.1f}% of the window's GPU time).\n")
    _, tab = table(ours)
    print(tab)
    marks = [i for i, r in enumerate(rows) if "argmax" in r[0].lower()]
    if len(marks) >= 2:
        step = rows[marks[-2] + 1: marks[-1] + 1]
        total, tab = table(step, top=16)
        print(f"\nOne decode step (launches between the last two argmax kernels): {len(step)} launches, {total / 1e3:.0f} us "
              f"(serialised, cold-cache timings).\n")
        print(tab)


if __name__ == "__main__":
    main()
