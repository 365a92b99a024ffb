"""CPU oracle for the MSDeformAttn hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under kinet_b200/ may import this package.  Allowed importers: tests/,
__graft_entry__.smoke(), and bench.py's cpu_baseline / --impl reference legs.
"""
from .msda_oracle import (  # noqa: F401
    build_oracle,
    oracle_forward,
    oracle_backward,
    oracle_max_threads,
    torch_port_forward,
)
