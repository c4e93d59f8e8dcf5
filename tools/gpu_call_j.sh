set -x
python - <<'PY'
import sys; sys.path.insert(0,'.')
from cuda_iblb_11_b200 import Params, Solver, FLAG_NO_GRAPH
for X,cn in ((576,12),):
    s = Solver(Params.channel(X,192,c_num=cn,c_space=48,c_fraction=cn, flags=FLAG_NO_GRAPH))
    s.step(200); s.sync(); s.close()
    s = Solver(Params.channel(X,192,c_num=cn,c_space=48,c_fraction=cn))
    s.step(200+64); s.sync(); s.close()
PY
