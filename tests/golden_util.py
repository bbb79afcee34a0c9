import glob
import os

import numpy as np
import scipy.sparse as sp

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    V = np.asfortranarray(z["V"])
    n = V.shape[0]
    A = sp.csc_matrix((z["aval"], z["rowidx"], z["colptr"]), shape=(n, n))
    levels, weighting = int(z["levels"]), int(z["weighting"])
    want = []
    for lvl in range(levels - 1):
        pre = f"L{lvl}_"
        want.append({k[len(pre):]: z[k] for k in z.files if k.startswith(pre)})
        want[-1]["radius"] = float(want[-1]["radius"])
    return V, A, levels, weighting, z["fine_val"], want
