"""Load the golden fixtures (outputs of the compiled reference, see tests/golden/make_golden.py)."""
import glob
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def case_names():
    return sorted(os.path.basename(p)[5:-4] for p in glob.glob(os.path.join(GOLDEN, "case_*.npz")))


def load_case(name):
    z = np.load(os.path.join(GOLDEN, "case_%s.npz" % name))
    c = {k: z[k] for k in z.files}
    flags = [str(f) for f in c["flags"]]
    c["flags"] = flags
    c["k"] = int(c["kmer_size"])
    c["rpb"] = int(c["reads_per_block"])
    c["abundance_set"] = "-abundance" in flags
    c["lossless"] = "-lossless" in flags
    c["has_qual"] = len(c["quals"]) > 0
    return c
