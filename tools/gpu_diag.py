#!/usr/bin/env python3
"""Stage-by-stage comparison of the CUDA path with the oracle / golden fixtures (development aid, GPU box)."""
import os
import sys
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from golden_util import case_names, load_case  # noqa: E402
from oracle import oracle_py as O  # noqa: E402
import leon_b200  # noqa: E402


def first_diff(a, b):
    n = min(len(a), len(b))
    d = np.nonzero(a[:n] != b[:n])[0]
    return int(d[0]) if len(d) else (n if len(a) != len(b) else -1)


def diag(name):
    c = load_case(name)
    ref = O.compress(c["bases"], c["offsets"], c["quals"] if c["has_qual"] else None, kmer_size=c["k"],
                     reads_per_block=c["rpb"], lossless=c["lossless"], abundance_set=c["abundance_set"])
    eng = leon_b200.Leon(kmer_size=c["k"], reads_per_block=c["rpb"], lossless=c["lossless"],
                         abundance=3 if c["abundance_set"] else None, has_qual=c["has_qual"])
    ok = True
    try:
        eng.upload(c["bases"], c["offsets"], c["quals"] if c["has_qual"] else None)
        cnt = eng.count()
        for key in ("cutoff", "nb_solids_cutoff", "nb_solid_total"):
            if cnt[key] != ref[key]:
                ok = False
                print("  COUNT %s: got %d want %d" % (key, cnt[key], ref[key]))
        if not np.array_equal(cnt["histogram"], ref["histogram"]):
            ok = False
            d = first_diff(cnt["histogram"], ref["histogram"])
            print("  COUNT histogram differs at bin %d: got %d want %d; nb_kmers=%d distinct=%d" % (
                d, cnt["histogram"][d], ref["histogram"][d], cnt["nb_kmers"], cnt["nb_distinct"]))
        bloom, bits = eng.bloom_build()
        if bits != ref["bloom_bits"] or not np.array_equal(bloom, ref["bloom"]):
            ok = False
            print("  BLOOM differs: bits %d/%d nbytes %d/%d first diff byte %d popcount %d/%d" % (
                bits, ref["bloom_bits"], len(bloom), len(ref["bloom"]), first_diff(bloom, ref["bloom"]),
                int(np.unpackbits(bloom).sum()), int(np.unpackbits(ref["bloom"]).sum())))
        r = eng.encode()
        if not np.array_equal(r["anchor_pos"], ref["anchor_pos"]):
            ok = False
            d = first_diff(r["anchor_pos"], ref["anchor_pos"])
            print("  ANCHOR pos differs first at read %d: got %d want %d (ndiff %d)" % (
                d, r["anchor_pos"][d], ref["anchor_pos"][d], int((r["anchor_pos"] != ref["anchor_pos"]).sum())))
        am = ref["anchor_pos"] >= 0
        if not np.array_equal(r["anchor_addr"][am], ref["anchor_addr"][am]):
            ok = False
            d = first_diff(r["anchor_addr"][am], ref["anchor_addr"][am])
            print("  ANCHOR addr differs first at anchored read #%d: got %d want %d" % (d, r["anchor_addr"][am][d], ref["anchor_addr"][am][d]))
        if r["anchor_count"] != ref["anchor_count"]:
            ok = False
            print("  ANCHOR count got %d want %d" % (r["anchor_count"], ref["anchor_count"]))
        ksz = np.uint64 if c["k"] <= 31 else np.uint64
        anchors = eng.debug_fetch("anchors", ksz)
        ra = ref["anchor_kmers"][:, 0] if c["k"] <= 31 else ref["anchor_kmers"].reshape(-1)
        if not np.array_equal(anchors, ra):
            ok = False
            print("  ANCHOR kmers differ at %d" % first_diff(anchors, ra))
        if not np.array_equal(r["anchors_dict"], ref["anchors_dict"]):
            ok = False
            print("  DICT stream differs: %d/%d bytes, first diff %d" % (len(r["anchors_dict"]), len(ref["anchors_dict"]),
                                                                      first_diff(r["anchors_dict"], ref["anchors_dict"])))
        if not np.array_equal(r["blocksizes"], ref["blocksizes"]):
            ok = False
            print("  BLOCKSIZES got %s want %s" % (r["blocksizes"][:10], ref["blocksizes"][:10]))
        if not np.array_equal(r["dna"], ref["dna"]):
            ok = False
            print("  DNA differs: %d/%d bytes, first diff %d" % (len(r["dna"]), len(ref["dna"]), first_diff(r["dna"], ref["dna"])))
        if c["has_qual"]:
            if not np.array_equal(r["qual_text"], ref["qual_text"]):
                ok = False
                d = first_diff(r["qual_text"], ref["qual_text"])
                print("  QUAL differs: %d/%d bytes, first diff %d (ndiff %d)" % (len(r["qual_text"]), len(ref["qual_text"]), d,
                      int((r["qual_text"][:len(ref["qual_text"])] != ref["qual_text"][:len(r["qual_text"])]).sum())))
            if not np.array_equal(r["qual_block_sizes"], ref["qual_block_sizes"]):
                ok = False
                print("  QUAL block sizes differ")
        # golden (reference) payloads as well
        for gk, rk in (("bloom", "bloom"), ("anchors_dict", "anchors_dict"), ("dna_concat", "dna"), ("blocksizes", "blocksizes")):
            if not np.array_equal(c[gk], r[rk]):
                ok = False
                print("  GOLDEN %s differs" % gk)
        print("%-18s %s  timings=%s" % (name, "OK" if ok else "MISMATCH", {k: round(v, 3) if isinstance(v, float) else v
                                                                          for k, v in eng.timings().items()}))
    except Exception:
        ok = False
        print("%-18s EXCEPTION" % name)
        traceback.print_exc()
    finally:
        eng.close()
    return ok


if __name__ == "__main__":
    names = sys.argv[1:] or case_names()
    res = [diag(n) for n in names]
    print("diag: %d/%d cases OK" % (sum(res), len(res)))
    sys.exit(0 if all(res) else 1)
