"""TEST INFRASTRUCTURE -- CPU oracles for the MSDeformAttn hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this package.  The product package never does.
"""
