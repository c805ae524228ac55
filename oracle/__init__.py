"""CPU oracle for the whisper-burn Q4_0 hot path.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package (whisper_burn_b200) never
imports this package.
"""
