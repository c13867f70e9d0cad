"""Experiment harness, system-metrics sidecar and analysis helpers (reference: ml/experiments/)."""
