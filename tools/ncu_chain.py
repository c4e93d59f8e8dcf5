"""Development (GPU box, under ncu --metrics gpu__time_duration.sum): a few iterations of the unfused
kernel chain on one grid, so that the per-kernel durations can be read from the launch list.
usage: python tools/ncu_chain.py X Y c_num c_space [flags]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_iblb_11_b200 import FLAG_NO_FUSE, FLAG_NO_GRAPH, Params, Solver

X, Y, cn, cs = [int(v) for v in sys.argv[1:5]]
flags = int(sys.argv[5]) if len(sys.argv) > 5 else (FLAG_NO_FUSE | FLAG_NO_GRAPH)
s = Solver(Params.channel(X, Y, c_num=cn, c_space=cs, c_fraction=cn, flags=flags))
s.step(40)
s.sync()
print("ok", s.last_step_seconds / 40 * 1e6, "us/step")
