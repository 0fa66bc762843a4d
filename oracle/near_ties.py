"""Near-tie probes for the sampling-MPC argmin (TEST INFRASTRUCTURE -- the checker, never the product).

north_star asks for an argmin index that is bit-exact against the NumPy oracle under fp64.  The kernel's costs agree
with the oracle's to ~1e-12 relative, not bit for bit (own sine, FMA-folded scaler, warm-started root), so the index is
exact wherever the best and second-best candidates are further apart than the two implementations' rounding noise.
This module PLANTS pairs of candidates whose ORACLE costs differ by a chosen relative gap (1e-8 ... 1e-15): the better
one at the HIGHER index and in another block / chunk than the worse one, so that a mis-ordered comparison and a
false tie (first index wins) both show up as a wrong index.  tests/test_gpu_parity.py::test_mpc_argmin_near_ties and
bench.py's ``parity`` leg run the CUDA step on these batches and record the smallest gap that still gives the
oracle's index.
"""
from __future__ import annotations

import numpy as np

from . import catenary_oracle as orc


def plant_near_ties(params, U, p_ref, gaps=(1e-8, 1e-10, 1e-12, 1e-13, 1e-14, 1e-15), h_perturb=0, axis=0, seed=0):
    """U (K,H,3) random candidates.  Returns a list of dict(U, want, gap, level, better) -- one batch per gap level
    and per sign ('better' = which of the pair is the oracle's winner: 'perturbed' or 'original').  The pair is the
    batch's best candidate and a copy of it with control [h_perturb, axis] moved by a few hundred ulps; `gap` is the
    ACTUAL relative gap of the two oracle costs in the returned batch."""
    rng = np.random.default_rng(seed)
    U = np.array(U, np.float64)
    K = U.shape[0]
    J = orc.mpc_rollout_costs(params, U, p_ref)
    b, Jb = orc.argmin_mpc(J)
    base = U[b].copy()
    # slope of the best candidate's cost along the perturbed control
    d0 = 1e-5 * max(1.0, abs(base[h_perturb, axis]))
    pm = np.stack([base, base])
    pm[0, h_perturb, axis] += d0
    pm[1, h_perturb, axis] -= d0
    Jp, Jm = orc.mpc_rollout_costs(params, pm, p_ref)
    slope = (Jp - Jm) / (2 * d0)
    if slope == 0.0:
        raise ValueError("the best candidate's cost does not depend on the perturbed control")
    out = []
    for level in gaps:
        for sign in (+1, -1):                       # +1: the perturbed copy is worse, -1: it is better
            target = sign * level * Jb
            centre = target / slope
            sweep = centre * (1.0 + 0.02 * np.arange(-40, 41))
            cand = np.repeat(base[None], sweep.size, axis=0)
            cand[:, h_perturb, axis] += sweep
            Jc = orc.mpc_rollout_costs(params, cand, p_ref)
            rel = (Jc - Jb) / Jb
            ok = (np.sign(rel) == sign) & (cand[:, h_perturb, axis] != base[h_perturb, axis])
            if not ok.any():
                continue                            # the gap is below one ulp of the cost for this candidate
            j = int(np.flatnonzero(ok)[np.argmin(np.abs(np.abs(rel[ok]) - level))])
            worse, better = (cand[j], base) if sign > 0 else (base, cand[j])
            # two slots in different 128-candidate blocks (and different eighths of the batch = host-pipeline chunks)
            i_lo = int(rng.integers(0, K // 2 - 1))
            i_hi = int(rng.integers(K // 2 + 1, K))
            if b in (i_lo, i_hi):
                i_lo = (i_lo + 1) % (K // 2 - 1)
            Ub = U.copy()
            Ub[b] = U[i_hi]                         # the original best row moves out of the way
            Ub[i_lo] = worse
            Ub[i_hi] = better
            Jn = orc.mpc_rollout_costs(params, Ub[[i_lo, i_hi]], p_ref)
            out.append(dict(U=Ub, want=i_hi, gap=float(abs(Jn[0] - Jn[1]) / abs(Jn[1])), level=float(level),
                            better="original" if sign > 0 else "perturbed", costs=(float(Jn[0]), float(Jn[1]))))
    return out


def margin(results):
    """results: list of (gap, index_matches).  Returns dict(min_gap_exact, first_mismatch_gap): the smallest tested
    gap down to which EVERY case at that gap or above matched, and the largest gap with a mismatch (None if none)."""
    res = sorted(results, key=lambda r: -r[0])
    min_ok, first_bad = None, None
    for gap, ok in res:
        if not ok:
            first_bad = gap
            break
        min_ok = gap
    return {"min_gap_exact": min_ok, "first_mismatch_gap": first_bad, "cases": len(res)}
