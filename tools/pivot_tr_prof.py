"""A few pivot count + transform launches at reduced row count (for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.argv = [sys.argv[0], "10000000", "50"]
exec(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "microbench_pivot.py")).read().split("# per-column timing")[0])
