"""Class shares of a step from an ncu launch list (gpu__time_duration.sum per launch, cold-cache and serialised) beside the
shares bench.py measured live with CUDA events in its timed region.

    python tools/launch_shares.py launches.csv bench_line.json

The launch list's absolute times are not bench values (ncu serialises the launches and runs them at its own clock); what has
to agree is each kernel class's SHARE of the step."""
from __future__ import annotations

import csv
import json
import sys
from collections import OrderedDict

CLASSES = OrderedDict([("logmel", ("logmel_kernel",)), ("normalize", ("normalize_apply_kernel",)), ("gemm_tcgen05", ("gemm_bf16_kernel",)),
                       ("attention_tcgen05", ("attention3_bf16_kernel", "attention_bf16_kernel")), ("layernorm", ("layernorm_kernel",)),
                       ("other", ("cast_bf16_kernel", "transpose_cast_kernel", "ln_fold_kernel"))])


def classify(name: str) -> str:
    for cls, keys in CLASSES.items():
        if any(k in name for k in keys):
            return cls
    return "unclassified:" + name.split("(")[0][-40:]


def main(csv_path: str, bench_path: str | None) -> None:
    rows = []
    with open(csv_path, newline="") as fh:
        lines = [ln for ln in fh if not ln.startswith("==")]
    reader = csv.DictReader(lines)
    for r in reader:
        if r.get("Metric Name") == "gpu__time_duration.sum":
            val = float(r["Metric Value"].replace(",", ""))
            unit = r.get("Metric Unit", "ns")
            ns = val * {"ns": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6, "nsecond": 1.0, "s": 1e9, "second": 1e9}.get(unit, 1.0)
            rows.append((r["Kernel Name"], ns))
    # one step = from one logmel launch sequence to the next: take the launches between the first and the second final LayerNorm
    ends = [i for i, (n, _) in enumerate(rows) if "layernorm_kernel" in n]
    if len(ends) >= 2:
        step = rows[ends[0] + 1: ends[1] + 1]
    else:
        step = rows
    tot = sum(ns for _, ns in step)
    by, cnt = OrderedDict(), OrderedDict()
    for name, ns in step:
        c = classify(name)
        by[c] = by.get(c, 0.0) + ns
        cnt[c] = cnt.get(c, 0) + 1
    bench = {}
    if bench_path:
        try:
            d = json.loads(open(bench_path).read().strip().splitlines()[-1])
            bench = {k: v["share_of_step"] for k, v in (d.get("rooflines") or {}).items()}
        except Exception as exc:  # the launch list stands on its own
            print(f"(bench line unreadable: {exc!r})")
    print(f"launches in {csv_path}: {len(rows)}; one step = {len(step)} launches, {tot / 1e6:.2f} ms under ncu (serialised, cold cache)")
    print(f"{'class':22s} {'launches':>8s} {'ncu ms':>9s} {'ncu share':>10s} {'bench share':>12s}")
    for c, ns in by.items():
        b = bench.get(c)
        print(f"{c:22s} {cnt[c]:8d} {ns / 1e6:9.3f} {ns / tot:10.4f} {('%12.4f' % b) if b is not None else '           -'}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None)
