"""Small solves of every kernel instantiation for compute-sanitizer (tools/sanitize.sh): force-feedback normal_1d and
point3d (with the friction cone), classical normal_1d, mixed free / contact instances, both line-search modes, a tick."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from ffmpc_cases import make_case  # noqa: E402
from franka_force_feedback_mpc_b200 import config, engine  # noqa: E402

B, N, iters = int(os.environ.get("SAN_B", 12)), int(os.environ.get("SAN_N", 12)), 3
for ff, nc, kw in ((True, 1, {}), (True, 3, dict(w_friction_cone=2.0e2, mu=0.3)), (False, 1, {}), (False, 3, {})):
    cfg, am, inp = make_case(ff, nc, B, N, 7, 0.6, **kw)
    for ls in ("queue", "spec"):
        eng = engine.FFMPCEngine(cfg, am, device=0, max_batch=B, line_search=ls, sub_batches=2)
        out = eng.solve_host(inp, iters, True)
        assert np.isfinite(out["cost"]).all()
        eng.close()
    print("ok", "ff" if ff else "classical", nc, flush=True)
print("done")
