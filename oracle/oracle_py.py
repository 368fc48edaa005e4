"""TEST INFRASTRUCTURE — ctypes binding of oracle/leon_oracle.c (CPU restatement of the reference path).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
The product package (leon_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")


def build(force=False):
    src = os.path.join(HERE, "leon_oracle.c")
    deps = [src, os.path.join(HERE, "leon_oracle.h"), os.path.join(HERE, "leon_tables.h")]
    if not force and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(d) for d in deps):
        return LIB
    subprocess.run(["gcc", "-O2", "-Wall", "-shared", "-fPIC", "-o", LIB, src], check=True, cwd=HERE)
    return LIB


class U128(C.Structure):
    _fields_ = [("lo", C.c_uint64), ("hi", C.c_uint64)]


class Params(C.Structure):
    _fields_ = [("kmer_size", C.c_int32), ("reads_per_block", C.c_int32), ("has_qual", C.c_int32),
                ("lossless", C.c_int32), ("abundance_set", C.c_int32)]


class Result(C.Structure):
    _fields_ = [
        ("cutoff", C.c_int32), ("nb_solids_cutoff", C.c_uint64), ("nb_solid_total", C.c_uint64),
        ("histogram", C.c_uint64 * 10001),
        ("bloom_bits", C.c_uint64), ("bloom_nbytes", C.c_uint64), ("bloom", C.POINTER(C.c_uint8)),
        ("anchor_count", C.c_uint64), ("anchors_dict_nbytes", C.c_uint64), ("anchors_dict", C.POINTER(C.c_uint8)),
        ("anchor_kmers", C.c_void_p),
        ("nblocks", C.c_uint64), ("blocksizes", C.POINTER(C.c_uint64)),
        ("dna_nbytes", C.c_uint64), ("dna", C.POINTER(C.c_uint8)),
        ("qual_nbytes", C.c_uint64), ("qual_text", C.POINTER(C.c_uint8)), ("qual_block_sizes", C.POINTER(C.c_uint64)),
        ("anchor_pos", C.POINTER(C.c_int32)), ("anchor_addr", C.POINTER(C.c_uint32)),
        ("read_count", C.c_uint64), ("total_dna_size", C.c_uint64),
        ("min_read_size", C.c_int32), ("max_read_size", C.c_int32),
        ("mc_total", C.c_uint64), ("mc_uniq_solid", C.c_uint64), ("mc_uniq_nosolid", C.c_uint64),
        ("mc_no_alternative", C.c_uint64), ("mc_multiple_solid", C.c_uint64), ("reads_without_anchor", C.c_uint64),
    ]


class Bloom(C.Structure):
    _fields_ = [("bits", C.POINTER(C.c_uint8)), ("nbytes", C.c_uint64), ("tai", C.c_uint64),
                ("reduced_tai", C.c_uint64), ("k", C.c_int), ("nhash", C.c_int)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        L.lo_hash64.restype = C.c_uint64
        L.lo_hash64.argtypes = [C.c_uint64, C.c_uint64]
        L.lo_bloom_seed.restype = C.c_uint64
        L.lo_bloom_seed.argtypes = [C.c_int]
        L.lo_rangecoder_kat.restype = C.c_uint64
        L.lo_rangecoder_kat.argtypes = [C.c_void_p, C.c_uint64]
        L.lo_compute_threshold.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_uint64)]
        L.lo_compress.restype = C.c_int
        L.lo_compress.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.POINTER(Params), C.POINTER(Result)]
        L.lo_free_result.argtypes = [C.POINTER(Result)]
        # __int128 by value follows the SysV ABI of a {lo, hi} pair of INTEGER class eightbytes
        L.lo_revcomp.restype = U128
        L.lo_revcomp.argtypes = [U128, C.c_int]
        L.lo_hash1.restype = C.c_uint64
        L.lo_hash1.argtypes = [U128, C.c_int, C.c_uint64]
        L.lo_simplehash16.restype = C.c_uint64
        L.lo_simplehash16.argtypes = [U128, C.c_int, C.c_int]
        L.lo_encode_kmer.restype = U128
        L.lo_encode_kmer.argtypes = [C.c_char_p, C.c_int]
        L.lo_bloom_init.argtypes = [C.POINTER(Bloom), C.c_uint64, C.c_int, C.c_int]
        L.lo_bloom_free.argtypes = [C.POINTER(Bloom)]
        L.lo_bloom_insert.argtypes = [C.POINTER(Bloom), U128]
        L.lo_bloom_contains.argtypes = [C.POINTER(Bloom), U128]
        L.lo_bloom_contains4.argtypes = [C.POINTER(Bloom), U128, C.c_int]
        _lib = L
    return _lib


def u128(v):
    return U128(v & 0xFFFFFFFFFFFFFFFF, v >> 64)


def from_u128(x):
    return (x.hi << 64) | x.lo


def _np(ptr, n, dtype):
    if n == 0 or not ptr:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(ptr, shape=(int(n),)).astype(dtype, copy=True)


def compress(bases, offsets, quals=None, kmer_size=31, reads_per_block=50000, lossless=False, abundance_set=False):
    """Run the whole restated path; returns a dict of numpy arrays / ints (copies; C memory freed)."""
    L = lib()
    bases = np.ascontiguousarray(bases, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.uint64)
    n = len(offsets) - 1
    has_qual = quals is not None and len(quals) > 0
    if has_qual:
        quals = np.ascontiguousarray(quals, np.uint8)
    p = Params(kmer_size, reads_per_block, int(has_qual), int(lossless), int(abundance_set))
    r = Result()
    rc = L.lo_compress(bases.ctypes.data, offsets.ctypes.data, n, quals.ctypes.data if has_qual else None,
                       C.byref(p), C.byref(r))
    if rc != 0:
        raise RuntimeError("lo_compress failed: %d" % rc)
    out = dict(
        cutoff=int(r.cutoff), nb_solids_cutoff=int(r.nb_solids_cutoff), nb_solid_total=int(r.nb_solid_total),
        histogram=np.array(r.histogram, np.uint64),
        bloom_bits=int(r.bloom_bits), bloom=_np(r.bloom, r.bloom_nbytes, np.uint8),
        anchor_count=int(r.anchor_count), anchors_dict=_np(r.anchors_dict, r.anchors_dict_nbytes, np.uint8),
        nblocks=int(r.nblocks), blocksizes=_np(r.blocksizes, 2 * r.nblocks, np.uint64),
        dna=_np(r.dna, r.dna_nbytes, np.uint8),
        qual_text=_np(r.qual_text, r.qual_nbytes, np.uint8),
        qual_block_sizes=_np(r.qual_block_sizes, r.nblocks, np.uint64) if has_qual else np.zeros(0, np.uint64),
        anchor_pos=_np(r.anchor_pos, n, np.int32), anchor_addr=_np(r.anchor_addr, n, np.uint32),
        read_count=int(r.read_count), total_dna_size=int(r.total_dna_size),
        min_read_size=int(r.min_read_size), max_read_size=int(r.max_read_size),
        mc_total=int(r.mc_total), reads_without_anchor=int(r.reads_without_anchor),
    )
    if r.anchor_count:
        raw = (C.c_uint64 * (2 * r.anchor_count)).from_address(r.anchor_kmers)
        out["anchor_kmers"] = np.array(raw, np.uint64).reshape(-1, 2)  # (lo, hi)
    else:
        out["anchor_kmers"] = np.zeros((0, 2), np.uint64)
    L.lo_free_result(C.byref(r))
    return out
