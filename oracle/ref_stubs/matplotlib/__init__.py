"""Stub for matplotlib (`mantrap/visualization/atomics.py:4-8`): plotting is out of scope, only the import must succeed.
Test infrastructure only (oracle/)."""
import sys
import unittest.mock

for _name in ("pyplot", "animation", "patches", "image", "offsetbox", "lines", "collections"):
    _mod = unittest.mock.MagicMock()
    sys.modules[f"matplotlib.{_name}"] = _mod
    globals()[_name] = _mod
