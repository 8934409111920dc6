"""CPU oracle: restatements of the reference's algorithms for the PC-sampling hot path.

TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this package; the product package (b_pinn_kalman_filter_b200) never does.
"""
