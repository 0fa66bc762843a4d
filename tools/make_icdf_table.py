#!/usr/bin/env python
"""Piecewise-cubic table of the inverse normal CDF used by the on-device control sampler (builder-defined; the
reference has no sampler -- SURVEY.md 8d config 3 asks for N(0, sigma^2) candidates reproducible in the oracle).

    z(r) = sign * y(r'),   r' = r if r < 2^31 else ~r,   sign = -1 if r < 2^31 else +1,   r' := max(r', 32)
    y(r') = -ndtri((r' + 0.5) / 2^32)  >= 0

y is tabulated on a floating-point-like grid: octave e = floor(log2 r') in 5..30, 32 sub-segments per octave taken from
the next five bits below the leading one, local variable t = (bits below those) / 2^(e-5) in [0, 1).  Each of the
26 x 32 = 832 segments carries a cubic c0 + t (c1 + t (c2 + t c3)) interpolating y at the four Chebyshev nodes of the
segment.  The table IS the definition of the sampler: the CUDA kernels (csrc/gen/icdf_table.cuh via codegen.py) and
oracle.sample_controls read the same numbers.  Writes artifacts/icdf_table.json.
"""
import json
import os
import sys

import numpy as np
from scipy.special import ndtri

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "..", "catenary-model-estimation-and-mpc-control-for-rov-tethered-systems_b200", "artifacts",
                   "icdf_table.json")
E_LO, E_HI, M = 5, 30, 5


def y_of(rp):
    return -ndtri((np.asarray(rp, np.float64) + 0.5) / 2.0 ** 32)


def main():
    deg = 3
    nodes = (np.cos(np.pi * (np.arange(deg + 1) + 0.5) / (deg + 1)) + 1) / 2
    V = np.vander(nodes, deg + 1, increasing=True)
    coef = np.zeros(((E_HI - E_LO + 1) << M, deg + 1))
    worst = 0.0
    worst_3 = 0.0
    for e in range(E_LO, E_HI + 1):
        w = 2.0 ** (e - M)                     # integers per segment
        for s in range(1 << M):
            lo = 2.0 ** e + s * w
            c = np.linalg.solve(V, y_of(lo + nodes * w))
            coef[((e - E_LO) << M) + s] = c
            tt = np.linspace(0.0, 1.0, 129)[:-1] if w >= 128 else np.arange(int(w)) / w
            yy = y_of(lo + tt * w)
            err = np.abs(((c[3] * tt + c[2]) * tt + c[1]) * tt + c[0] - yy)
            worst = max(worst, float(err.max()))
            if (yy <= 3.3).any():
                worst_3 = max(worst_3, float(err[yy <= 3.3].max()))
    out = {"e_lo": E_LO, "e_hi": E_HI, "sub_bits": M, "degree": deg, "max_abs_err": worst,
           "max_abs_err_within_3p3_sigma": worst_3,
           "definition": __doc__.strip().split("\n\n")[1],
           "coef": [[float(x) for x in row] for row in coef]}
    with open(OUT, "w") as fh:
        json.dump(out, fh)
    print(f"{coef.shape[0]} segments, max |err| {worst:.3e} (|z| <= 3.3: {worst_3:.3e}) -> {os.path.relpath(OUT)}")


if __name__ == "__main__":
    sys.exit(main())
