"""CPU oracle for the OCSMesh raster hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``ocsmesh_b200/`` may import this package.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs use it, as the checker or as the timed CPU baseline.
"""
from .ocsmesh_oracle import *  # noqa: F401,F403
