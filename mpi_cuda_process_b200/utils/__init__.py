"""Timing, clocks, metrics, checkpointing and logging (all absent in the reference)."""
from . import checkpoint, clocks, log, metrics, timing  # noqa: F401
