"""Quick end-to-end check used for the first GPU call (python -m tests.gpu_quick)."""
import time
import numpy as np
from rayne_b200 import capi, scenegen as sg
from tests.helpers import check_against_oracle, standard_cameras

t = time.time()
for kw in (dict(n_trees=3, branching=4, depth=8, seed=2), dict(ragged_nodes=60000, depth=8, seed=5, random_bounds=True)):
    got, ref = check_against_oracle(sg.make_scene(**kw), standard_cameras())
    print("parity ok", kw, got["stats"], [len(l) for l in got["lists"]], flush=True)
print("elapsed", time.time() - t)
