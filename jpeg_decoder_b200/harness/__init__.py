"""Harness: the bundled baseline-JPEG encoder and synthetic picture generator (inputs only)."""
from .synth import synth_jpeg, synth_rgb, encode_rgb, build_harness, HARNESS_SO  # noqa: F401
