"""Timing, clock sampling and misc helpers."""
