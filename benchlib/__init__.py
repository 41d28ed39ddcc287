"""Pieces of bench.py (the driver's contract lives in bench.py itself): workloads and synthetic data, clock sampling,
device measurements, the multi-GPU layouts, and -- in cpu.py only -- the legs that may execute oracle/."""
