"""CPU oracle for the SRBD-MPC hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package.  The product (g1_locomotion_b200) never does; it fails loudly without its CUDA library.
"""
