"""Parity bar shared by the GPU (and emulator) tests.

north_star: per-node gains K/k, the cost and the gaps within 1e-8 relative at a fixed iteration count.  The checker is
the fp64 oracle (oracle/libffmpc_oracle.so).  Ten FDDP iterations amplify rounding differently in two independent fp64
implementations, so an instance that misses 1e-8 against the fp64 oracle is not judged by that oracle's own noise but by
an ARBITER: the same oracle source built in IEEE binary128 (oracle/libffmpc_oracle_quad.so).  Rule, per output array:

    max_i err(engine vs quad)  <=  max(1e-8, 3 * max_i err(fp64 oracle vs quad))      over the instances sent to the arbiter

i.e. the engine may not be further from the binary128 result than the fp64 restatement of the reference algorithm itself.
`iters` / `ok` (the accept / reject decisions) must agree with the fp64 oracle exactly.
"""
from __future__ import annotations

import os

import numpy as np

TOL = 1e-8
KEYS = ("K", "k", "cost", "xs", "us")


def inst_rel(a, b):
    """per-instance max-abs error relative to that instance's largest reference entry"""
    a = np.asarray(a, float).reshape(a.shape[0], -1)
    b = np.asarray(b, float).reshape(b.shape[0], -1)
    return np.abs(a - b).max(axis=1) / np.maximum(np.abs(b).max(axis=1), 1e-300)


def _errs(out, ref, key):
    """per-instance error of `out[key]` against `ref[key]` on the scale the quantity is used on"""
    B = ref["cost"].shape[0]
    if key == "gaps":  # gaps are ~0 on a closed trajectory: absolute, on the scale of the state
        return np.abs(out[key] - ref[key]).reshape(B, -1).max(axis=1) / np.maximum(1.0, np.abs(ref[key]).reshape(B, -1).max(axis=1))
    e = inst_rel(out[key], ref[key])
    if key == "k":
        # k is the feed-forward INCREMENT of us and tends to 0 as the solve converges: its error is measured on the scale
        # of the controls it corrects (relative to |k| itself it is 0/0 at convergence)
        sc = np.abs(ref["k"]).reshape(B, -1).max(axis=1) / np.maximum(np.abs(ref["us"]).reshape(B, -1).max(axis=1), 1e-300)
        e = e * np.minimum(sc, 1.0)
    return e


def check_vs_oracle(out, cfg, am, inp, iters, fixed, orc, keys=KEYS + ("gaps",), tol=TOL, ref=None, label="", arbiter="lazy"):
    """Asserts the parity bar for one batched solve; returns {key: (max err vs fp64 oracle, n instances sent to the arbiter)}.
    arbiter = "lazy": binary128 only for the instances that miss `tol` against the fp64 oracle; "always": for the whole batch."""
    nthreads = os.cpu_count() or 1
    if ref is None:
        ref = orc.solve_batch(cfg, am, inp, iters, fixed, nthreads=nthreads)
    assert np.array_equal(out["iters"], ref["iters"]), (label, "iters")
    assert np.array_equal(out["ok"], ref["ok"]), (label, "ok")
    report = {}
    e_all = {k: _errs(out, ref, k) for k in keys}
    need = np.zeros(inp.B, dtype=bool)
    for k in keys:
        need |= ~(e_all[k] <= tol)
    if arbiter == "always":
        need[:] = True
    idx = np.nonzero(need)[0]
    if idx.size:
        sub = inp.slice(idx)
        quad = orc.solve_batch(cfg, am, sub, iters, fixed, nthreads=nthreads, _use=orc.quad_lib())
        out_s = {k: out[k][idx] for k in set(keys) | {"us", "cost"}}
        ref_s = {k: ref[k][idx] for k in set(keys) | {"us", "cost"}}
        for k in keys:
            eg, eo = _errs(out_s, quad, k), _errs(ref_s, quad, k)
            assert eg.max() <= max(tol, 3.0 * eo.max()), (label, k, "engine vs binary128", float(eg.max()), "fp64 oracle vs binary128", float(eo.max()),
                                                           "instances", idx[:8].tolist())
    for k in keys:
        report[k] = (float(e_all[k].max()), int(idx.size))
    # contact force left by the last rollout (knots 0 / 1): same bar on contact instances, NaN in free mode
    if "f_pred" in out and "f_pred" in ref:
        c = inp.mode == 1
        if c.any():
            ef = inst_rel(out["f_pred"][c], ref["f_pred"][c])
            bad = ~(ef <= tol)
            assert not (bad & ~need[c]).any(), (label, "f_pred", float(ef.max()))
        assert np.isnan(out["f_pred"][~c]).all(), (label, "f_pred free mode")
    return report
