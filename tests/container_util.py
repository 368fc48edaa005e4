"""Helpers of the `.leon` container tests: decode a container with the compiled reference (oracle/_ref/leon -d, built
from /root/reference by oracle/build_ref.sh; it travels to the GPU box with the snapshot) and walk the HDF5 structures
of a file with a small independent parser (no HDF5 library in this image)."""
import os
import re
import struct
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")


def ref_leon(kmer_size):
    """Path of the compiled reference binary able to handle kmer_size, or None when it was not built."""
    p = os.path.join(REF_DIR, "leon64" if kmer_size > 31 else "leon")
    return p if os.path.exists(p) else None


def result_from_golden(c):
    """The datasets the reference stored for a golden case, shaped like Leon.compress()'s result."""
    r = dict(bloom=c["bloom"], bloom_bits=int(c["bloom_bits"]), anchors_dict=c["anchors_dict"],
             anchor_count=int(c["anchor_count"]), blocksizes=c["blocksizes"], dna=c["dna_concat"],
             read_count=int(c["meta_readcount"]), total_dna_size=int(c["meta_totalDnaSize"]),
             min_read_size=int(c["meta_minSequenceSize"]), max_read_size=int(c["meta_maxSequenceSize"]))
    if c["has_qual"]:
        r.update(qual_text=c["qual_text_concat"], qual_block_sizes=c["qual_text_sizes"])
    return r


def reference_decode(leon_file, kmer_size, fastq):
    """Runs the reference decoder on leon_file; returns (reads, quals) as lists of bytes (quals None for FASTA)."""
    wd = os.path.dirname(leon_file)
    res = subprocess.run([ref_leon(kmer_size), "-file", leon_file, "-d", "-nb-cores", "1"], cwd=wd,
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and "EXCEPTION" not in res.stdout + res.stderr, res.stdout[-2000:] + res.stderr[-2000:]
    dec = leon_file[:-len(".leon")] + ".d"
    # header-less mode: the decoder numbers the reads and prints one stray numbered line at the end of every full
    # block (tests/golden/make_golden.py), so the numbered pseudo-headers are dropped rather than counted
    hdr = re.compile(rb"^[>@] \d+$")
    lines = [l.rstrip(b"\n") for l in open(dec, "rb")]
    lines = [l for l in lines if not hdr.match(l)]
    if not fastq:
        return lines, None
    return lines[0::3], lines[2::3]


def reads_of(bases, offsets):
    o = np.asarray(offsets).astype(np.int64)
    return [bases[o[i]:o[i + 1]].tobytes() for i in range(len(o) - 1)]


# ---- a minimal reader of the structures leon_write_container emits (HDF5 1.8 format, version-0 superblock) ----------

class H5Mini:
    """Parses superblock v0, v1 object headers, symbol-table groups (B-tree v1 + SNOD + local heap), contiguous
    datasets and variable-length string attributes (global heap).  Raises AssertionError on any inconsistency."""

    def __init__(self, path):
        self.b = open(path, "rb").read()
        b = self.b
        assert b[:8] == b"\x89HDF\r\n\x1a\n" and b[8] == 0 and b[13] == 8 and b[14] == 8
        self.leaf_k, self.int_k = struct.unpack_from("<HH", b, 16)
        base, _, eof, _ = struct.unpack_from("<QQQQ", b, 24)
        assert base == 0 and eof == len(b), (eof, len(b))
        name_off, ohdr, cache, _, btree, heap = struct.unpack_from("<QQIIQQ", b, 56)
        assert cache == 1
        self.root = self._group(ohdr)
        assert (self.root["_btree"], self.root["_heap"]) == (btree, heap)

    def _messages(self, at):
        b = self.b
        ver, _, nmsgs, nlink, size = struct.unpack_from("<BBHII", b, at)
        assert ver == 1 and nlink == 1 and at % 8 == 0
        p, end, out = at + 16, at + 16 + size, []
        while p < end:
            t, n, fl = struct.unpack_from("<HHB", b, p)
            assert n % 8 == 0
            out.append((t, b[p + 8:p + 8 + n]))
            p += 8 + n
        assert p == end and len(out) == nmsgs
        return out

    def _heap_name(self, heap, off):
        b = self.b
        assert b[heap:heap + 4] == b"HEAP"
        size, free, data = struct.unpack_from("<QQQ", b, heap + 8)
        assert free == 1 and off < size
        e = b.index(b"\0", data + off)
        return b[data + off:e].decode()

    def _tree_leaves(self, node, heap, lo_key):
        """Yields SNOD addresses in key order, checking the keys that bracket every child."""
        b = self.b
        assert b[node:node + 4] == b"TREE" and b[node + 4] == 0
        level, used = b[node + 5], struct.unpack_from("<H", b, node + 6)[0]
        assert used <= 2 * self.int_k
        p = node + 24
        key = struct.unpack_from("<Q", b, p)[0]
        assert self._heap_name(heap, key) == lo_key
        p += 8
        for _ in range(used):
            child, nkey = struct.unpack_from("<QQ", b, p)
            p += 16
            prev, hi = self._heap_name(heap, key), self._heap_name(heap, nkey)
            if level:
                yield from self._tree_leaves(child, heap, prev)
            else:
                yield child, prev, hi
            key = nkey

    def _group(self, ohdr):
        msgs = self._messages(ohdr)
        assert len(msgs) == 1 and msgs[0][0] == 0x11
        btree, heap = struct.unpack("<QQ", msgs[0][1][:16])
        b, out, names = self.b, {"_btree": btree, "_heap": heap}, []
        for snod, lo, hi in self._tree_leaves(btree, heap, ""):
            assert b[snod:snod + 4] == b"SNOD" and b[snod + 4] == 1
            n = struct.unpack_from("<H", b, snod + 6)[0]
            assert 0 < n <= 2 * self.leaf_k
            for i in range(n):
                noff, oh, cache = struct.unpack_from("<QQI", b, snod + 8 + 40 * i)
                name = self._heap_name(heap, noff)
                assert lo < name <= hi
                names.append(name)
                out[name] = self._group(oh) if cache == 1 else self._dataset(oh)
        assert names == sorted(names) and len(set(names)) == len(names)
        return out

    def _vlen(self, elem):
        b = self.b
        n, col, idx = struct.unpack("<IQI", elem)
        assert b[col:col + 4] == b"GCOL" and b[col + 4] == 1
        size = struct.unpack_from("<Q", b, col + 8)[0]
        assert size >= 4096 and col + size <= len(b)
        p = col + 16
        while p + 16 <= col + size:
            i, refs, _, osz = struct.unpack_from("<HHIQ", b, p)
            if i == 0:
                assert p + osz == col + size
                break
            if i == idx:
                assert osz == n
                return b[p + 16:p + 16 + n].decode()
            p += 16 + ((osz + 7) & ~7)
        raise AssertionError("global heap object %d not found" % idx)

    def _dataset(self, ohdr):
        b, d = self.b, {"attrs": {}}
        for t, body in self._messages(ohdr):
            if t == 1:
                assert body[0] == 1 and body[1] == 1
                d["n"] = struct.unpack_from("<Q", body, 8)[0]
            elif t == 3:
                assert body[:12] == bytes([0x10, 0, 0, 0, 1, 0, 0, 0, 0, 0, 8, 0])
            elif t == 8:
                assert body[0] == 3 and body[1] == 1
                d["addr"], d["size"] = struct.unpack_from("<QQ", body, 2)
            elif t == 0xC:
                assert body[0] == 1
                nl, dl, sl = struct.unpack_from("<HHH", body, 2)
                p = 8
                name = body[p:p + nl - 1].decode()
                p += (nl + 7) & ~7
                assert body[p] == 0x19 and body[p + 1] == 1
                p += (dl + 7) & ~7
                assert body[p + 1] == 1 and struct.unpack_from("<Q", body, p + 8)[0] == 1
                p += (sl + 7) & ~7
                d["attrs"][name] = self._vlen(body[p:p + 16])
            else:
                raise AssertionError("unexpected message %#x" % t)
        assert d["n"] == d["size"]
        d["data"] = np.frombuffer(b, np.uint8, d["n"], d["addr"]) if d["n"] else np.zeros(0, np.uint8)
        return d


def ref_h5dump():
    p = os.path.join(REF_DIR, "gatb-h5dump")
    return p if os.path.exists(p) else None


def ref_dataset(leon_file, dset, scratch):
    """Bytes of one dataset of an HDF5 file, read by the reference's own gatb-h5dump."""
    out = os.path.join(scratch, "dump.bin")
    if os.path.exists(out):
        os.remove(out)
    subprocess.run([ref_h5dump(), "-d", dset, "-b", "LE", "-o", out, leon_file], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return np.fromfile(out, dtype=np.uint8)


def ref_attr(leon_file, attr):
    txt = subprocess.run([ref_h5dump(), "-a", attr, leon_file], check=True, capture_output=True, text=True).stdout
    return re.search(r'\(0\): "(.*)"', txt).group(1)
