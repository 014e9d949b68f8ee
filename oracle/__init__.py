"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

CPU restatement of the reference hot path.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this package; graspqp_b200/ never does.
"""
