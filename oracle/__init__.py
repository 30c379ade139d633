"""CPU oracles — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference`
legs may import this package.  The product (4paradigm-k8s-device-plugin_b200/) never does.
"""
