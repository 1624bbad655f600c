"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's paged-attention hot path.

Nothing in the product package imports this directory.  Allowed importers: tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
"""
