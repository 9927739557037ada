"""one 256x256x64 solve with bench.py's knobs (the bottom depth is the headline's 256x256x4): target for ncu captures"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
from common import Problem
from somar_b200 import capi
pb = Problem(n=(256, 256, 64), map_name="stretched", aniso=16.0, parts=(1, 8, 1))
lev = pb.level(capi, relaxMode=1, **bench.SOLVE)
st = lev.solve(lev.field(0), lev.field(0, pb.rhs()))
print(st["iters"], st["exit_status"], st["bottom_iters_total"], st["solve_ms"])
