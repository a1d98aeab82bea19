"""Generates tests/golden/*.json from the REFERENCE ITSELF: /root/reference/src/phase1.c,
linear.c, fixed.c, config.c compiled unmodified into oracle/_ref (see oracle/Makefile),
run on the reference's own example instance and on small seeded synthetic instances.
Run in the build container (needs /root/reference):  python tests/golden/make_golden.py
The GPU box has no /root/reference; tests there read only the committed JSON.
"""
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
README_IN = "/root/reference/examples/readme_example.in"
README_BETA = [0.984331027786964, 0.792399824970372, 0.754117840176144, 0.592849130685193, 0.057351715952213]  # README.md:87

CONFIGS = [  # (bits_p1, bits_p2, precision, precision_p2)
    (64, 64, 56, 0),
    (32, 32, 24, 0),
    (64, 32, 56, 24),
]


def hexrows(a):
    return [["%016x" % int(v) for v in row] for row in a]


def main():
    po.build(ref=True)
    out = {"ti_key": po.TI_KEY.hex(), "ot_key": po.OT_KEY.hex(), "readme_beta": README_BETA, "cases": []}
    X, y, cs = po.read_instance(README_IN)
    instances = [("readme_example", README_IN, X, y, cs)]
    tmp = tempfile.mkdtemp()
    for name, n, d, starts, seed in [("syn_96x7", 96, 7, [0, 2, 5], 11), ("syn_33x6_odd_n", 33, 6, [0, 3], 12),
                                     ("syn_64x9_p4", 64, 9, [0, 2, 4, 6], 13)]:
        Xs, ys, _ = po.generate_lin_regression(n, d, seed=seed)
        path = os.path.join(tmp, name + ".in")
        po.write_instance(path, Xs, ys, starts)
        instances.append((name, path, Xs, ys, starts))
    for name, path, Xi, yi, starts in instances:
        n, d = Xi.shape
        for b1, b2, prec, prec2 in CONFIGS:
            R = po.RefLib(b1, b2)
            Xe, ye, _ = R.read_inputs(path, n, d, prec)
            for use_ot in (False, True):
                shares, _, sent = R.run(path, d, len(starts), prec, prec2, use_ot)
                out["cases"].append({
                    "instance": name, "n": n, "d": d, "col_start": list(starts),
                    "bits_p1": b1, "bits_p2": b2, "precision": prec, "precision_p2": prec2, "use_ot": use_ot,
                    "X": [repr(float(v)) for v in Xi.ravel()], "y": [repr(float(v)) for v in yi],
                    "encoded_X_first": int(Xe[0, 0]), "encoded_y_first": int(ye[0]),
                    "shares": hexrows(shares),
                    "bytes_sent": [[int(v) for v in row] for row in sent],
                })
    with open(os.path.join(HERE, "phase1_ref_shares.json"), "w") as f:
        json.dump(out, f, indent=0)
    print("wrote", len(out["cases"]), "cases")


if __name__ == "__main__":
    main()
