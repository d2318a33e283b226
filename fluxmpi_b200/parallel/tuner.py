"""Per-message-size algorithm and launch-shape selection from measured bus bandwidth.

SURVEY §5.1 asks for the variant (one-shot / two-shot P2P / NVLS) to be "selected per message
size from a measured busbw table". ``bench/tune.py`` measures every candidate on the actual
box; ``bench/make_tuned.py`` keeps the fastest per (op, world, size) and writes
``fluxmpi_b200/parallel/tuned.json`` (override with ``$FMPI_TUNE_FILE``). At run time a
collective looks up the entry whose size is nearest on a log scale; without an entry the C
side's built-in heuristics (``FMPI_ALGO_AUTO``) decide.

The reference has nothing comparable: every call is whatever libmpi does.
"""
from __future__ import annotations

import ctypes as C
import json
import math
import os
from pathlib import Path
from typing import Optional

from .. import _native
from .._native import Launch

_ALGO = {"auto": _native.ALGO_AUTO, "oneshot": _native.ALGO_ONESHOT, "twoshot": _native.ALGO_TWOSHOT,
         "nvls": _native.ALGO_NVLS, "pipe": _native.ALGO_PIPE, "pull_p2p": _native.ALGO_PULL_P2P}

# Sizes from which the library's own choice (FMPI_ALGO_AUTO) is a kernel the round-1 grids never
# measured: the pull + re-emit broadcast on symmetric memory (csrc/kernels/collectives.cu,
# bcast_mode(): 268 / 371 / 470 / 491 GB/s at 16 MB / 64 MB / 256 MB / 1 GB on 8 GPUs against
# 134-394 for the tuned root-multicast rows). A table without "pipe" rows must not shadow it.
_AUTO_FROM = {"bcast_symm": 8 << 20}


class Tuner:
    def __init__(self, world: int, multicast: bool, path: Optional[str] = None):
        self.world = world
        self.multicast = multicast
        self.table: dict[str, list[tuple[int, str, int, int]]] = {}
        self._cache: dict[tuple[str, int], object] = {}
        p = Path(path or os.environ.get("FMPI_TUNE_FILE") or Path(__file__).with_name("tuned.json"))
        if p.exists() and os.environ.get("FMPI_NO_TUNE", "0") in ("0", ""):
            try:
                data = json.loads(p.read_text())
            except (OSError, ValueError):
                data = {}
            for op, rows in data.get(str(world), {}).items():
                ok = [(int(r["bytes"]), r["algo"], int(r["blocks"]), int(r["threads"])) for r in rows
                      if r["algo"] != "nvls" or multicast]
                if ok:
                    self.table[op] = sorted(ok)
        self.auto_from = {op: lim for op, lim in _AUTO_FROM.items()
                          if not any(r[1] == "pipe" for r in self.table.get(op, ()))}

    def choice(self, op: str, nbytes: int):
        """(algo name, blocks, threads) or None."""
        rows = self.table.get(op)
        if not rows or nbytes <= 0 or nbytes * 2 < rows[0][0]:
            return None          # below the measured range: the library's latency defaults apply
        if nbytes >= self.auto_from.get(op, 1 << 62):
            return None          # the library's large-message kernels (see _AUTO_FROM)
        lg = math.log2(nbytes)
        best = min(rows, key=lambda r: abs(math.log2(r[0]) - lg))
        return best[1], best[2], best[3]

    def launch(self, op: str, nbytes: int):
        """ctypes byref(Launch) for this call, or None to let the library decide. Cached per
        (op, power-of-two size bucket): the eager collectives are host-latency sensitive."""
        key = (op, max(nbytes, 1).bit_length())
        hit = self._cache.get(key, 0)
        if hit != 0:
            return hit
        ch = self.choice(op, nbytes)
        if ch is None:
            ref = None
        else:
            obj = Launch(ch[1], ch[2], _ALGO[ch[0]])
            ref = (C.byref(obj), obj)          # keep the struct alive next to its reference
        self._cache[key] = ref
        return ref
