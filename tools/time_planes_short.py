"""cfg4 / cfg2-disc / cfg3-disc lines of tools/time_planes.py only (A/B of kernel variants)."""
import os, sys, runpy
sys.argv = [sys.argv[0]] + sys.argv[1:]
os.environ["TAC_PLANES_SHORT"] = "1"
runpy.run_path(os.path.join(os.path.dirname(os.path.abspath(__file__)), "time_planes.py"), run_name="__main__")
