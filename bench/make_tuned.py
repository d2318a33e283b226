#!/usr/bin/env python
"""Fold bench/tune.py outputs into fluxmpi_b200/parallel/tuned.json.

    python bench/make_tuned.py gpurun_out/tune_2.json gpurun_out/tune_4.json gpurun_out/tune_8.json

Keeps, per (world, op, message size), the candidate with the lowest device time; existing
entries for worlds that are not in the inputs are preserved.
"""
from __future__ import annotations

import json
import sys
from pathlib import Path

OPS = ["allreduce_any", "reduce_any", "bcast_any", "allreduce_symm", "bcast_symm"]
OUT = Path(__file__).resolve().parent.parent / "fluxmpi_b200" / "parallel" / "tuned.json"


def main(paths):
    table = json.loads(OUT.read_text()) if OUT.exists() else {}
    for p in paths:
        d = json.loads(Path(p).read_text())
        w = str(d["world"])
        per_op = {}
        for op in OPS:
            best = {}
            for r in d["rows"]:
                k = op + "_us"
                if k not in r:
                    continue
                cur = best.get(r["bytes"])
                if cur is None or r[k] < cur[k]:
                    best[r["bytes"]] = r
            per_op[op] = [{"bytes": b, "algo": r["algo"], "blocks": r["blocks"], "threads": r["threads"],
                           "us": r[op + "_us"], "busbw_gbs": r[op + "_busbw"]} for b, r in sorted(best.items())]
        table[w] = per_op
    table["_comment"] = ("fastest (algo, blocks, threads) per op / world / message size, measured by "
                         "bench/tune.py on the gpurun B200 boxes (CUDA events, max over ranks)")
    OUT.write_text(json.dumps(table, indent=1))
    print("wrote", OUT)


if __name__ == "__main__":
    main(sys.argv[1:])
