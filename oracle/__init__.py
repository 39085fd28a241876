"""TEST INFRASTRUCTURE ONLY: CPU oracle for the Event-3DGS splatting hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package. The product path (event_3dgs_b200/) never does.
"""
