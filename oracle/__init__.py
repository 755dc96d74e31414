"""CPU oracle (TEST INFRASTRUCTURE): ctypes binding of oracle/rbd_oracle.c.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this.
"""
from .oracle import Oracle, build_oracle  # noqa: F401
