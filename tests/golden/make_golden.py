"""Regenerates the committed golden fixtures.

rootsolve_pdf_p16.json -- transcribed BY HAND from the reference's frozen vignette output
  (/root/reference/inst/doc/rootSolve.pdf p.16; model and call in
  /root/reference/vignettes/rootSolve.Rnw:686-739).  These are the reference's own numbers; this
  script only rewrites the file so that the provenance is in one place.
structure_hashes.json  -- sha256 of ian/jan/igp/jgp for a set of grids, produced by the pinned CPU
  oracle (oracle/stodes_oracle.c).  The reference holds no element-wise fixture for these arrays
  (SURVEY.md 8(c)); the closed-form counts it does imply are stored alongside.

Run from the repository root:  python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

PDF_P16 = {
    "source": "rootSolve 1.8.2.4 inst/doc/rootSolve.pdf p.16; vignettes/rootSolve.Rnw:686-739",
    "n": 10, "y0": 10.0, "tol": 1e-10, "lrw_given": 100000,
    "parms": [1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 0.025, 1.0],
    "nnz": 6400, "ngp": 12, "lrw_used": 82366,
    "nnz_estimated_by_R": 6910, "function_calls_estimated_by_R": 1001,
    "precis": [1.105000e+01, 3.917674e-07, 2.751792e-16],
    "steady": True,
}

# rootsolve_pdf_p8_anox.json -- transcribed by hand from inst/doc/rootSolve.pdf p.8-9: the 4-variable anoxic
# sediment model (vignettes/rootSolve.Rnw:360-403) solved by the dense sibling `stode(y, fun, parms, pos=TRUE)`.
# Same Newton skeleton (errset, Perturb, clamp, the two tests); stodes reaches the same root.
PDF_P8_ANOX = {
    "source": "rootSolve 1.8.2.4 inst/doc/rootSolve.pdf p.8-9; vignettes/rootSolve.Rnw:360-403",
    "parms": [1.0, 100.0, 0.1, 1.0, 1.0, 1.0, 100.0, 10000.0, 0.0],
    "parm_names": ["D", "Flux", "r", "rox", "ks", "ks2", "BO2", "BSO4", "BHS"],
    "y0": [1.0, 1.0, 1.0, 1.0],
    "y": [1000.012783, 6.825178, 9996.587411, 3.412589],
    "SumS": 10000.0,
    "precis": [2.549712e+03, 5.753884e+01, 2.039705e+01, 8.527476e+00, 2.168616e+00, 1.515096e-01, 7.266703e-04,
               1.664189e-08],
    "steady": True,
}

# rootsolve_man_steady2d.json -- the arguments the two steady.2D man-page examples are CALLED with
# (man/steady.2D.Rd:132-143 and :193-206).  The second lrw is exactly the work length the path needs for that
# example (the oracle fails with one double less), i.e. a value the author read back from attributes()$dims.
MAN_STEADY2D = {
    "source": "rootSolve 1.8.2.4 man/steady.2D.Rd:106-207",
    "example1": {"dimens": [100, 100], "cyclicBnd": None, "lrw_given": 1000000, "y0": 10.0, "tol": 1e-10,
                 "parms": {"Dx": 1, "Dy": 1, "dx": 1, "dy": 1, "r": 0.025, "BND": 1}},
    "example2": {"dimens": [50, 100], "cyclicBnd": 2, "lrw_given": 283800, "lrw_needed": 283800, "y0": 10.0, "tol": 1e-10,
                 "parms": {"Dx": 1, "Dy": 1, "dx": 1, "dy": 1, "r": 0.025, "BND": 1, "extra_flux": 10, "exchange": 10}},
}

GRIDS = [
    ("1D", 1, [50], [0]), ("1D", 5, [1000], [0]), ("1D", 2, [30], [1]),
    ("2D", 1, [50, 50], [0, 0]), ("2D", 1, [100, 100], [0, 0]), ("2D", 2, [17, 9], [0, 0]), ("2D", 1, [50, 100], [0, 1]),
    ("3D", 1, [10, 10, 10], [0, 0, 0]), ("3D", 1, [16, 16, 16], [0, 0, 0]), ("3D", 2, [8, 8, 8], [0, 0, 0]),
    ("3D", 1, [6, 6, 6], [1, 1, 1]), ("3D", 1, [32, 32, 32], [0, 0, 0]),
]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.int32).tobytes()).hexdigest()


def main():
    from oracle import oracle as O
    json.dump(PDF_P16, open(os.path.join(HERE, "rootsolve_pdf_p16.json"), "w"), indent=1)
    json.dump(PDF_P8_ANOX, open(os.path.join(HERE, "rootsolve_pdf_p8_anox.json"), "w"), indent=1)
    json.dump(MAN_STEADY2D, open(os.path.join(HERE, "rootsolve_man_steady2d.json"), "w"), indent=1)
    out = []
    for kind, nspec, dims, cyc in GRIDS:
        ian, jan = O.sparse_structure(kind, nspec, dims, cyc)
        ngp, igp, jgp = O.jgroup(ian, jan)
        out.append(dict(kind=kind, nspec=nspec, dims=dims, cyclic=cyc, n=len(ian) - 1, nnz=len(jan), ngp=ngp,
                        group_sizes=np.diff(igp).tolist(), ian=sha(ian), jan=sha(jan), igp=sha(igp), jgp=sha(jgp)))
    json.dump(out, open(os.path.join(HERE, "structure_hashes.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
