#!/usr/bin/env python
"""Regression gate over the benchmark's machine-readable results.

`p2p_matrix --json FILE` (engine) and `oracle/_ref/p2p_matrix_nccl --json FILE` (the reference's
NCCL path, same loops) write one JSON record per measured cell.  This script pairs the records
of two such files by (mode, timing, src, dst, bytes) and fails — exit code 1 — when the engine
falls below the thresholds:

    python scripts/regression_gate.py ENGINE.json BASELINE.json [--min-ratio 1.0]
                                      [--min-link-frac 0.75 --from-bytes 67108864]

  --min-ratio R       bandwidth cells: engine GB/s >= R x baseline; latency cells: engine us <= baseline / R
  --min-link-frac F   uni-directional cells of at least --from-bytes must reach F x 900 GB/s per direction
  --allow N           tolerate up to N failing cells (noise)

The reference keeps its results in a git-ignored result.txt (reference .gitignore:1-2) and has
no notion of a threshold; this is the gate a CI job on a B200 box would run.
"""
from __future__ import annotations

import argparse
import json
import sys

LINK_GBPS = 900.0
LOWER_IS_BETTER = {"latency"}


def load(path):
    out = {}
    with open(path) as f:
        for line in f:
            line = line.strip()
            if not line:
                continue
            r = json.loads(line)
            if r.get("mode") == "verify":
                continue
            key = (r["mode"], r.get("timing", "dev"), r.get("src", -1), r.get("dst", -1), r["bytes"])
            out[key] = r
    return out


def gate(engine, baseline, min_ratio=1.0, min_link_frac=None, from_bytes=64 << 20):
    """Returns (rows, failures); a row is (key, engine value, baseline value, ratio, verdict)."""
    rows, failures = [], []
    for key in sorted(set(engine) & set(baseline)):
        e, b = engine[key], baseline[key]
        mode = key[0]
        if mode in LOWER_IS_BETTER:
            ev, bv = e["us"], b["us"]
            ratio = bv / ev if ev > 0 else float("inf")   # > 1 means the engine is faster
            unit = "us"
        else:
            ev, bv = e["GBps"], b["GBps"]
            ratio = ev / bv if bv > 0 else float("inf")
            unit = "GB/s"
        ok = ratio >= min_ratio
        why = "" if ok else f"ratio {ratio:.3f} < {min_ratio}"
        if ok and min_link_frac is not None and mode in ("uni", "stream") and key[4] >= from_bytes:
            frac = ev / LINK_GBPS
            if frac < min_link_frac:
                ok, why = False, f"{frac:.3f} of link < {min_link_frac}"
        rows.append((key, ev, bv, unit, ratio, ok, why))
        if not ok:
            failures.append((key, why))
    return rows, failures


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("engine")
    ap.add_argument("baseline")
    ap.add_argument("--min-ratio", type=float, default=1.0)
    ap.add_argument("--min-link-frac", type=float, default=None)
    ap.add_argument("--from-bytes", type=int, default=64 << 20)
    ap.add_argument("--allow", type=int, default=0)
    ap.add_argument("--quiet", action="store_true")
    args = ap.parse_args(argv)
    engine, baseline = load(args.engine), load(args.baseline)
    rows, failures = gate(engine, baseline, args.min_ratio, args.min_link_frac, args.from_bytes)
    if not rows:
        print("no cell is present in both files", file=sys.stderr)
        return 2
    if not args.quiet:
        print(f"{'mode':9s} {'timing':6s} {'src':>3s} {'dst':>3s} {'bytes':>11s} {'engine':>10s} {'baseline':>10s} {'unit':5s} {'ratio':>7s}")
        for (mode, timing, src, dst, nbytes), ev, bv, unit, ratio, ok, why in rows:
            print(f"{mode:9s} {timing:6s} {src:3d} {dst:3d} {nbytes:11d} {ev:10.2f} {bv:10.2f} {unit:5s} {ratio:7.3f} {'' if ok else 'FAIL ' + why}")
    worst = min(r[4] for r in rows)
    print(f"{len(rows)} cells compared, {len(failures)} below threshold, worst ratio {worst:.3f}")
    return 1 if len(failures) > args.allow else 0


if __name__ == "__main__":
    sys.exit(main())
