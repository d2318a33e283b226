"""Preferences, checkpointing and tracing helpers."""
from . import prefs  # noqa: F401
