"""CPU-only checks of the boundary: the C-ABI library loads and exports every symbol include/leon_b200.h declares,
fails loudly without a device, and the synthetic generator is deterministic."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _built():
    import __graft_entry__ as g
    g.build()


def test_abi_exports_every_declared_symbol():
    _built()
    import leon_b200
    L = leon_b200.load_library()
    hdr = open(os.path.join(ROOT, "include", "leon_b200.h")).read()
    declared = set(re.findall(r"\b(leon_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"leon_opts", "leon_reads"}
    assert declared, "no declarations found"
    for sym in sorted(declared):
        assert hasattr(L, sym), "libleon_b200.so does not export %s" % sym
    assert set(leon_b200.ABI_SYMBOLS) == declared
    assert L.leon_b200_abi_version() == 3


def test_no_cpu_fallback():
    """Without a CUDA device leon_create must fail (LEON_E_CUDA); with one this test is skipped."""
    _built()
    import leon_b200
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    with pytest.raises(leon_b200.LeonError) as e:
        leon_b200.Leon()
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_product_never_touches_oracle():
    for d, _, files in os.walk(os.path.join(ROOT, "leon_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(d, f)).read()
                assert "oracle" not in txt.replace("oracle/", "").lower() or f == "kmer_math.cuh", f
                assert "oracle_py" not in txt and "liboracle" not in txt and "leon_oracle" not in txt, f


def test_synth_is_deterministic():
    from leon_b200 import synth
    g = synth.make_genome(5000, 1)
    a = synth.sample_reads(g, 300, 100, 2, with_quals=True, n_rate=0.001)
    b = synth.sample_reads(g, 300, 100, 2, with_quals=True, n_rate=0.001)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[2], b[2])
    assert set(np.unique(a[0]).tolist()) <= set(b"ACGTN")
    assert a[2].min() >= 33 + 2 and a[2].max() <= 33 + 40
