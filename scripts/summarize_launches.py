"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list of `bench.py`: kernels of the LAST
ELBO+grad evaluation (from one sep_apply_C_kernel / build_C_kernel launch of the fused closure to the next),
grouped by kernel, with each kernel's share of the evaluation.

usage: python scripts/summarize_launches.py gpurun_out/launches.csv [launches_per_eval [first_launch]] > profiles/rN_launches_summary.txt
"""
import csv
import re
import sys
from collections import OrderedDict


def short(name):
    name = re.sub(r"\(anonymous namespace\)::|sgp::|void ", "", name)
    m = re.match(r"([A-Za-z_0-9:<>]+?)(<.*?>)?\(", name)
    if not m:
        return name[:60]
    targs = m.group(2) or ""
    targs = re.sub(r"\((int|bool)\)", "", targs)
    return m.group(1) + targs


def main():
    path = sys.argv[1]
    rows = []
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for rec in csv.DictReader(lines):
        if rec["Metric Name"] == "gpu__time_duration.sum":
            rows.append((rec["Kernel Name"], float(rec["Metric Value"].replace(",", "")) * 1e-6))
    # one launch per evaluation marks its start: sep_apply_C_kernel on the separable path (C itself is built later, at the
    # head of the inducing-point chain), build_C_kernel otherwise
    marker = "sep_apply_C_kernel" if any("sep_apply_C_kernel" in k for k, _ in rows) else "build_C_kernel"
    starts = [i for i, (k, _) in enumerate(rows) if marker in k]
    if len(starts) < 2:
        raise SystemExit("need at least two evaluations in the launch list")
    per_eval = int(sys.argv[2]) if len(sys.argv) > 2 else starts[-1] - starts[-2]
    # the last FULL evaluation of the device-resident timed loop: a window of per_eval launches from a build_C
    cand = [s for s in starts if s + per_eval <= len(rows)]
    s0 = int(sys.argv[3]) if len(sys.argv) > 3 else cand[-1]
    ev = rows[s0:s0 + per_eval]
    total = sum(t for _, t in ev)
    agg = OrderedDict()
    for k, t in ev:
        a = agg.setdefault(short(k), [0.0, 0])
        a[0] += t
        a[1] += 1
    print(f"# {path}: launches {s0}..{s0 + per_eval - 1} = one ELBO+grad evaluation, {per_eval} launches, "
          f"{total:.3f} ms (serialised, cold-cache per-launch times: shares, not absolutes, carry over)")
    for k, (t, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"{t:9.3f} ms {100 * t / total:5.1f}%  x{n:3d}  {k}")


if __name__ == "__main__":
    main()
