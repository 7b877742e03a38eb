"""ORACLE — CPU/fp32 PyTorch restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under multimodal-transformer-rl_b200/ may import this package; only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs do, and only as the checker / the timed CPU baseline.
"""
