"""CPU oracle for libexamg -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  The product path (exastencils_b200/) must never do so.
"""
