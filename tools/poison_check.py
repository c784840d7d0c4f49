import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ate_replication_causalml_b200.engine import Engine
from ate_replication_causalml_b200 import synth
d = synth.make_dataset(3000, 9)
eng = Engine(0)
fid = synth.fold_ids(3000, 10, 7)
xw = np.column_stack([d["x_raw"], d["w"]])
try:
    g = eng.cv_glmnet(xw, d["y"], "gaussian", penalty_factor=np.r_[np.ones(9), 0.0], foldid=fid)
    print("gauss cv ok", g["lmu"], g["idx_1se"])
except Exception as e:
    print("FAIL", str(e)[:400])
