"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name.

    python profiles/summarize_launches.py gpurun_out/launches.csv
Per-launch times under ncu are cold-cache and serialised: read the SHARES, not the absolutes.
"""
import collections
import csv
import re
import sys


def load(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    return [(r["Kernel Name"], float(r["Metric Value"].replace(",", "")), r["Grid Size"], r["Block Size"]) for r in rows]


def short(n):
    n = re.sub(r"b200llm::\(anonymous namespace\)::|void |unnamed>::", "", n)
    n = re.sub(r"\(.*", "", n)
    return n[:70]


def table(rows, title):
    agg = collections.OrderedDict()
    for n, v, g, b in rows:
        a = agg.setdefault(short(n), [0, 0.0, g])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"## {title}: {len(rows)} launches, {tot / 1000:.1f} us total")
    print("| kernel | launches | total us | avg us | share | grid |")
    print("|---|---:|---:|---:|---:|---|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| {k} | {a[0]} | {a[1] / 1000:.1f} | {a[1] / a[0] / 1000:.2f} | {100 * a[1] / tot:.1f}% | {a[2]} |")
    print()


if __name__ == "__main__":
    rows = load(sys.argv[1])
    names = [r[0] for r in rows]
    idx = [i for i, n in enumerate(names) if "prefill_attn" in n]
    if idx:
        first_pf = min(i for i, n in enumerate(names) if "prefill_meta" in n or "init_slots" in n)
        end_pf = next(i for i in range(max(idx), len(names)) if "select_kernel" in names[i]) + 1
        table(rows[first_pf:end_pf], "prefill")
        table(rows[end_pf:], "decode steps")
    else:
        table(rows, "all")
