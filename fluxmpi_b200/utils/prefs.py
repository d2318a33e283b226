"""Persisted preferences (the analogue of the reference's Preferences.jl use).

Reference: ``/root/reference/src/FluxMPI.jl:16-31`` reads ``FluxMPIDisableCUDAMPISupport`` from
``LocalPreferences.toml`` at module load and ``:51-56`` writes it; a restart is needed for the
change to apply. Here the switch is "disable the NVLink peer-to-peer / multicast kernels and
fall back to NCCL (device tensors) — the A/B baseline path".

File: ``$FMPI_PREFS_FILE`` or ``./LocalPreferences.toml``; section ``[FluxMPIB200]``.
"""
from __future__ import annotations

import os
import tomllib
from pathlib import Path

SECTION = "FluxMPIB200"
KEY_DISABLE_P2P = "FluxMPIDisableP2PSupport"


def prefs_path() -> Path:
    return Path(os.environ.get("FMPI_PREFS_FILE", "LocalPreferences.toml"))


def load_all() -> dict:
    p = prefs_path()
    if not p.exists():
        return {}
    try:
        with p.open("rb") as f:
            return tomllib.load(f)
    except (tomllib.TOMLDecodeError, OSError):
        return {}


def load_preference(key: str, default=None):
    return load_all().get(SECTION, {}).get(key, default)


def _fmt(v) -> str:
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, (int, float)):
        return repr(v)
    return '"' + str(v).replace("\\", "\\\\").replace('"', '\\"') + '"'


def set_preference(key: str, value) -> Path:
    data = load_all()
    data.setdefault(SECTION, {})[key] = value
    lines = []
    for sec, kv in data.items():
        if not isinstance(kv, dict):
            continue
        lines.append(f"[{sec}]")
        for k, v in kv.items():
            if isinstance(v, dict):
                continue
            lines.append(f"{k} = {_fmt(v)}")
        lines.append("")
    p = prefs_path()
    p.write_text("\n".join(lines))
    return p
