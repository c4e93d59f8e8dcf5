"""Development tool (GPU box): launch each of the small kernels beside the LB kernel a few times so
that one `ncu --set full` run can capture them (tools/gpu_call.sh stage ncu_markers):
ib_kinematics_batch, ib_interp, ib_spread (unfused chain), ib_tile_lists (fused path), lb_init_t
(iblb_set_fields) and lb_moments_t (iblb_get_fields), on the cilia12.sh grid (C3) and, for the two
host-buffer kernels, on a 4096x4096 channel where they move enough bytes to be bandwidth-bound."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

from cuda_iblb_11_b200 import FLAG_NO_FUSE, FLAG_NO_GRAPH, Params, Solver

kw = dict(c_num=12, c_space=48, c_fraction=12)
s = Solver(Params.channel(576, 192, flags=FLAG_NO_FUSE | FLAG_NO_GRAPH, **kw))
s.step(3)      # ib_kinematics_batch x1, ib_interp x3, ib_spread x3
s.sync()
s.close()
s = Solver(Params.channel(576, 192, flags=FLAG_NO_GRAPH, **kw))
s.step(2)      # ib_kinematics_batch, ib_tile_lists, lb_ib_tile_step x2
s.sync()
s.close()
s = Solver(Params.channel(4096, 4096, gx=1e-7))
rho = np.ones(s.size)
u = np.zeros(2 * s.size)
s.set_fields(rho, u)   # lb_init_t (one chunk: 3 x 16.7 M doubles = 403 MB > one half of the stage)
s.step(2)
s.fields(rho, u)       # lb_moments_t
s.close()
print("done")
