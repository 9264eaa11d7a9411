"""CPU oracle of the Shifu GBT/RF training hot path -- TEST INFRASTRUCTURE ONLY (see shifu_oracle.cpp).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
