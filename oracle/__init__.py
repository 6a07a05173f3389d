"""CPU oracle for the rasterizer hot path -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product (``eags-slam_b200/``) never does.

PARITY UNPINNED: see the header of ``eags_oracle.c`` and DESIGN.md section 3.
"""
from .oracle import (  # noqa: F401
    OracleSettings,
    build,
    forward,
    backward,
    num_threads,
    set_num_threads,
    lib_path,
)
