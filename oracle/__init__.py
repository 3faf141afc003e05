"""CPU oracle for the visual_slam hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under visual_slam_b200/ may import this package; only tests/, bench.py's
cpu_baseline / --impl reference legs and __graft_entry__.smoke() use it, as the checker.
"""
