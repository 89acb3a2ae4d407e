"""Stub for cyipopt 0.1.9 (`mantrap/solver/ipopt.py:3`), which is not installable here (no network, no gfortran).
Test infrastructure only (oracle/)."""


def problem(*args, **kwargs):
    raise ImportError("cyipopt is unavailable in this container")
