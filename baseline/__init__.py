"""Baseline harness: runs the unmodified reference (installed under ``baseline/_ref``)."""
