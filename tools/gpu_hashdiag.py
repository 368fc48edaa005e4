import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from golden_util import load_case
from oracle import oracle_py as O
import leon_b200
import ctypes as C

for name in ("fa_k63", "fq_k47_lossy", "fa_k31_blocks"):
    c = load_case(name); k = c["k"]
    eng = leon_b200.Leon(kmer_size=k, reads_per_block=c["rpb"])
    eng.upload(c["bases"], c["offsets"]); eng.count(); bloom, bits = eng.bloom_build()
    rng = np.random.default_rng(5)
    n = 4000
    vals = [int.from_bytes(rng.bytes(16), "little") & ((1 << (2 * k)) - 1) for _ in range(n)]
    km = np.array([[v & (2**64 - 1), v >> 64] for v in vals], np.uint64)
    out = eng.debug_hashes(km)
    L = O.lib()
    names = ["canon_lo", "canon_hi", "hp_lo", "hp_hi", "racine", "prefval"] + ["sh%d" % i for i in range(1, 7)]
    bad = {nm: 0 for nm in names}
    cano2 = [0, 1, 2, 3, 4, 5, 3, 7, 8, 9, 0, 4, 9, 13, 1, 5]
    ex = None
    for i, v in enumerate(vals):
        rc = O.from_u128(L.lo_revcomp(O.u128(v), k)); cn = min(v, rc)
        hp = (cn >> 2) & ((1 << (2 * (k - 2))) - 1); hr = O.from_u128(L.lo_revcomp(O.u128(hp), k - 2)); hp = min(hp, hr)
        rac = L.lo_hash1(O.u128(hp), k, L.lo_bloom_seed(0)) % bits
        pv = cano2[((cn >> (2 * (k - 1))) & 3) * 4 + (cn & 3)]
        want = [cn & (2**64 - 1), cn >> 64, hp & (2**64 - 1), hp >> 64, rac, pv] + [L.lo_simplehash16(O.u128(hp), k, j) for j in range(1, 7)]
        for j, nm in enumerate(names):
            if int(out[i, j]) != want[j]:
                bad[nm] += 1
                if ex is None:
                    ex = (hex(v), nm, hex(int(out[i, j])), hex(want[j]))
    print(name, "k=%d" % k, bad, "first:", ex)
    eng.close()
