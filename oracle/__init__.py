"""CPU oracle for the CHRONOS hot path.  TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs
only -- never by the product package (chronos_b200), which must fail loudly without
its CUDA library rather than fall back to anything here.
"""
