"""CPU oracle for the routing hot path — TEST INFRASTRUCTURE ONLY.

Importable from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
Nothing under routest_b200/ may import this package.
"""
