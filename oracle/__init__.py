"""CPU oracle for the CRAM slice-codec hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  The product (cramtools_b200) never does.
"""
