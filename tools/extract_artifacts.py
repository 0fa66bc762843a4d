#!/usr/bin/env python
"""Extract the model artefacts the hot path needs from the reference checkout.

Runs ONLY in the build container (needs /root/reference). Writes
<package>/artifacts/model.json, which is committed: the GPU box has no
/root/reference, and build-time code generation (codegen.py) reads this file.

Sources (reference file:line):
  saved_models/equations_dtheta_dt.csv  (all rows; shipped row = complexity 13, line 10)
  saved_models/equations_dgamma_dt.csv  (all rows; shipped row = complexity 3, line 3)
  saved_models/eq_dtheta_dt.txt:1 / eq_dgamma_dt.txt:1  (which complexity was shipped)
  saved_models/scaler.pkl               (StandardScaler mean_/scale_, simple.py:57-58,114)
"""
import csv
import json
import os
import re
import sys
import warnings

REF = os.environ.get("CAT_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(HERE, "..", "catenary-model-estimation-and-mpc-control-for-rov-tethered-systems_b200")


def rows(path):
    with open(path, newline="") as fh:
        return [
            {"complexity": int(r["complexity"]), "loss": float(r["loss"]),
             "equation": r["equation"], "sympy_format": r["sympy_format"]}
            for r in csv.DictReader(fh)
        ]


def shipped_complexity(txt_path):
    # eq_*.txt is a truncated pandas repr; its first line is "complexity   <n>"
    with open(txt_path) as fh:
        m = re.match(r"\s*complexity\s+(\d+)", fh.readline())
    return int(m.group(1))


def main():
    import joblib
    warnings.filterwarnings("ignore")
    sm = os.path.join(REF, "saved_models")
    scaler = joblib.load(os.path.join(sm, "scaler.pkl"))
    out = {
        "source": "eather0056/Catenary-Model-Estimation-and-MPC-Control-for-ROV-Tethered-Systems saved_models/",
        "n_features": int(scaler.n_features_in_),
        "scaler_mean": [float(v) for v in scaler.mean_],
        "scaler_scale": [float(v) for v in scaler.scale_],
        "dtheta_dt": {"shipped_complexity": shipped_complexity(os.path.join(sm, "eq_dtheta_dt.txt")),
                      "rows": rows(os.path.join(sm, "equations_dtheta_dt.csv"))},
        "dgamma_dt": {"shipped_complexity": shipped_complexity(os.path.join(sm, "eq_dgamma_dt.txt")),
                      "rows": rows(os.path.join(sm, "equations_dgamma_dt.csv"))},
    }
    dst = os.path.join(PKG, "artifacts", "model.json")
    with open(dst, "w") as fh:
        json.dump(out, fh, indent=1)
    print("wrote", os.path.normpath(dst))


if __name__ == "__main__":
    sys.exit(main())
