"""ctypes binding of include/leon_b200.h (no torch types, no compute on the Python side)."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libleon_b200.so")

ABI_SYMBOLS = [
    "leon_b200_abi_version", "leon_create", "leon_destroy", "leon_last_error", "leon_alloc_pinned", "leon_free_pinned",
    "leon_compress", "leon_upload", "leon_count_resident", "leon_count_begin", "leon_count", "leon_count_finish",
    "leon_encode_block", "leon_block_wait", "leon_encode_finish", "leon_stream_prefetch", "leon_bloom_build", "leon_encode", "leon_run_device", "leon_fetch",
    "leon_get_timings", "leon_get_kernel_times", "leon_set_stream", "leon_comm_unique_id", "leon_comm_init",
    "leon_comm_batch_reads", "leon_debug_bloom_query", "leon_debug_hashes", "leon_debug_fetch", "leon_write_container",
    "leon_headers_encode", "leon_headers_free", "leon_compress_file", "leon_decompress_file", "leon_read_file", "leon_file_reads_free",
    "leon_decode_dna", "leon_decode_dna_device", "leon_decompress_file_device", "leon_decoded_block", "leon_decoded_free", "leon_headers_decode",
    "leon_kmer_offsets_noalloc", "leon_container_open", "leon_container_close", "leon_container_has", "leon_container_dataset", "leon_container_attr",
]


class LeonError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("leon_b200 error %d: %s" % (code, msg))
        self.code = code


class Opts(C.Structure):
    _fields_ = [("kmer_size", C.c_int32), ("reads_per_block", C.c_int32), ("has_qual", C.c_int32), ("lossless", C.c_int32),
                ("abundance_set", C.c_int32), ("device", C.c_int32), ("anchor_batch", C.c_int32), ("dict_on_device", C.c_int32),
                ("qual_zlib_on_device", C.c_int32), ("skip_qual_text", C.c_int32), ("skip_read_arrays", C.c_int32),
                ("reserved", C.c_int32 * 5)]


class Reads(C.Structure):
    _fields_ = [("bases", C.c_void_p), ("offsets", C.c_void_p), ("quals", C.c_void_p), ("nreads", C.c_uint64)]


class CountResult(C.Structure):
    _fields_ = [("cutoff", C.c_int32), ("nb_solids_cutoff", C.c_uint64), ("nb_solid_total", C.c_uint64),
                ("nb_distinct", C.c_uint64), ("nb_kmers", C.c_uint64), ("histogram", C.POINTER(C.c_uint64))]


class Result(C.Structure):
    _fields_ = [
        ("count", CountResult),
        ("bloom", C.POINTER(C.c_uint8)), ("bloom_nbytes", C.c_uint64), ("bloom_bits", C.c_uint64),
        ("anchors_dict", C.POINTER(C.c_uint8)), ("anchors_dict_nbytes", C.c_uint64), ("anchor_count", C.c_uint64),
        ("nblocks", C.c_uint64), ("blocksizes", C.POINTER(C.c_uint64)),
        ("dna", C.POINTER(C.c_uint8)), ("dna_nbytes", C.c_uint64),
        ("qual_text", C.POINTER(C.c_uint8)), ("qual_nbytes", C.c_uint64), ("qual_block_sizes", C.POINTER(C.c_uint64)),
        ("read_count", C.c_uint64), ("total_dna_size", C.c_uint64),
        ("min_read_size", C.c_int32), ("max_read_size", C.c_int32),
        ("anchor_pos", C.POINTER(C.c_int32)), ("anchor_addr", C.POINTER(C.c_uint32)),
        ("reads_without_anchor", C.c_uint64),
        ("qual_z", C.POINTER(C.c_uint8)), ("qual_z_nbytes", C.c_uint64), ("qual_z_block_sizes", C.POINTER(C.c_uint64)),
    ]


class StreamPlan(C.Structure):
    _fields_ = [("total_reads", C.c_uint64), ("total_bases", C.c_uint64), ("count_passes", C.c_uint32), ("reserved", C.c_uint32 * 3)]


class BlockOut(C.Structure):
    _fields_ = [("first_block", C.c_uint64), ("nblocks", C.c_uint64), ("read_count", C.c_uint64),
                ("blocksizes", C.POINTER(C.c_uint64)), ("dna", C.POINTER(C.c_uint8)), ("dna_nbytes", C.c_uint64),
                ("qual_text", C.POINTER(C.c_uint8)), ("qual_nbytes", C.c_uint64), ("qual_block_sizes", C.POINTER(C.c_uint64)),
                ("qual_z", C.POINTER(C.c_uint8)), ("qual_z_nbytes", C.c_uint64), ("qual_z_block_sizes", C.POINTER(C.c_uint64)),
                ("anchor_pos", C.POINTER(C.c_int32)), ("anchor_addr", C.POINTER(C.c_uint32)), ("reserved", C.c_uint64 * 4)]


class Timings(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("h2d_ms", "count_ms", "sort_ms", "bloom_ms", "anchor_ms", "walk_ms", "qual_ms",
                                         "entropy_ms", "d2h_ms", "total_ms")] + \
               [("kernel_launches", C.c_uint64), ("sort_passes", C.c_uint32), ("anchor_batches", C.c_uint32),
                ("anchor_rounds", C.c_uint32), ("sort_pass_ms", C.c_float), ("sort_keys", C.c_uint64),
                ("dict_host_ms", C.c_float), ("dict_wait_ms", C.c_float), ("dict_copy_wait_ms", C.c_float),
                ("reserved", C.c_float * 3), ("nsym_total", C.c_uint64)]


class ContainerInfo(C.Structure):
    _fields_ = [("kmer_size", C.c_int32), ("is_fasta", C.c_int32), ("has_header", C.c_int32), ("reserved0", C.c_int32),
                ("first_header", C.c_void_p), ("first_header_nbytes", C.c_uint64), ("header_nblocks", C.c_uint64),
                ("header_blocksizes", C.c_void_p), ("header_data", C.c_void_p), ("reserved", C.c_uint64 * 4)]


class HeaderStream(C.Structure):
    _fields_ = [("first_header", C.POINTER(C.c_uint8)), ("first_header_nbytes", C.c_uint64), ("nblocks", C.c_uint64),
                ("blocksizes", C.POINTER(C.c_uint64)), ("data", C.POINTER(C.c_uint8)), ("data_nbytes", C.c_uint64),
                ("coded_serially", C.c_int32), ("reserved", C.c_int32)]


class FileOpts(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("kmer_size", "reads_per_block", "lossless", "abundance_set", "noheader", "noqual",
                                         "seq_only", "device", "nthreads", "stream_reads", "count_passes")] + [("reserved", C.c_int32 * 5)]


class FileStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("nreads", "total_bases", "nblocks", "input_bytes", "output_bytes",
                                          "header_bytes_in", "header_bytes_out", "dna_bytes_out")] + \
               [(n, C.c_int32) for n in ("is_fasta", "has_header", "headers_coded_serially", "reserved0")] + \
               [(n, C.c_float) for n in ("parse_ms", "compress_ms", "header_ms", "write_ms", "total_ms")] + \
               [("reserved1", C.c_float * 3), ("gpu", Timings), ("cutoff", C.c_int32), ("abundance_auto", C.c_int32)] + \
               [(n, C.c_uint64) for n in ("nb_solid_kmers", "anchor_count", "reads_without_anchor", "bloom_bytes", "anchors_dict_bytes",
                                          "seq_bytes_out", "qual_bytes_out")]


class FileReads(C.Structure):
    _fields_ = [("reads", Reads), ("headers", C.POINTER(C.c_uint8)), ("header_offsets", C.POINTER(C.c_uint64)),
                ("is_fastq", C.c_int32), ("threads_used", C.c_int32)]


class DnaStream(C.Structure):
    _fields_ = [("kmer_size", C.c_int32), ("max_read_size", C.c_int32),
                ("bloom", C.c_void_p), ("bloom_nbytes", C.c_uint64), ("bloom_bits", C.c_uint64),
                ("anchors_dict", C.c_void_p), ("anchors_dict_nbytes", C.c_uint64), ("anchor_count", C.c_uint64),
                ("nblocks", C.c_uint64), ("blocksizes", C.c_void_p), ("dna", C.c_void_p), ("dna_nbytes", C.c_uint64)]


class KernelTime(C.Structure):
    _fields_ = [("name", C.c_char * 32), ("ms_total", C.c_float), ("launches", C.c_uint32)]


_lib = None


def load_library(path=LIB_PATH):
    """Loads libleon_b200.so; raises if it was not built (no fallback of any kind)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise LeonError(-2, "%s not built: run `python -c 'import __graft_entry__ as g; g.build()'`" % path)
    L = C.CDLL(path)
    L.leon_create.argtypes = [C.POINTER(Opts), C.POINTER(C.c_void_p)]
    L.leon_destroy.argtypes = [C.c_void_p]
    L.leon_last_error.restype = C.c_char_p
    L.leon_last_error.argtypes = [C.c_void_p]
    L.leon_alloc_pinned.restype = C.c_void_p
    L.leon_alloc_pinned.argtypes = [C.c_size_t]
    L.leon_free_pinned.argtypes = [C.c_void_p]
    L.leon_compress.argtypes = [C.c_void_p, C.POINTER(Reads), C.POINTER(Result)]
    L.leon_upload.argtypes = [C.c_void_p, C.POINTER(Reads)]
    L.leon_count_resident.argtypes = [C.c_void_p, C.POINTER(CountResult)]
    L.leon_count_begin.argtypes = [C.c_void_p, C.POINTER(StreamPlan), C.POINTER(C.c_uint32)]
    L.leon_count.argtypes = [C.c_void_p, C.POINTER(Reads), C.c_int]
    L.leon_count_finish.argtypes = [C.c_void_p, C.POINTER(CountResult)]
    L.leon_encode_block.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(Reads), C.POINTER(BlockOut)]
    L.leon_stream_prefetch.argtypes = [C.c_void_p, C.POINTER(Reads), C.c_int]
    L.leon_block_wait.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(BlockOut)]
    L.leon_encode_finish.argtypes = [C.c_void_p, C.POINTER(Result)]
    L.leon_bloom_build.argtypes = [C.c_void_p, C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.leon_encode.argtypes = [C.c_void_p, C.POINTER(Result)]
    L.leon_run_device.argtypes = [C.c_void_p]
    L.leon_fetch.argtypes = [C.c_void_p, C.POINTER(Result)]
    L.leon_get_timings.argtypes = [C.c_void_p, C.POINTER(Timings)]
    L.leon_get_kernel_times.argtypes = [C.c_void_p, C.POINTER(KernelTime), C.c_int, C.POINTER(C.c_int)]
    L.leon_set_stream.argtypes = [C.c_void_p, C.c_void_p]
    L.leon_comm_unique_id.argtypes = [C.c_void_p]
    L.leon_comm_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    L.leon_comm_batch_reads.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
    L.leon_debug_bloom_query.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]
    L.leon_debug_hashes.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    L.leon_kmer_offsets_noalloc.argtypes = [C.c_int32, C.c_void_p, C.c_uint64, C.c_int32, C.c_void_p, C.c_void_p]
    L.leon_debug_fetch.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
    L.leon_write_container.argtypes = [C.c_char_p, C.POINTER(Result), C.POINTER(ContainerInfo)]
    L.leon_headers_encode.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int32, C.c_int32, C.POINTER(C.POINTER(HeaderStream))]
    L.leon_headers_free.argtypes = [C.POINTER(HeaderStream)]
    L.leon_headers_free.restype = None
    L.leon_compress_file.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(FileOpts), C.POINTER(FileStats)]
    L.leon_decompress_file.argtypes = [C.c_char_p, C.c_char_p, C.c_int32, C.POINTER(FileStats)]
    L.leon_read_file.argtypes = [C.c_char_p, C.c_int, C.c_int, C.POINTER(C.POINTER(FileReads))]
    L.leon_file_reads_free.argtypes = [C.POINTER(FileReads)]
    L.leon_file_reads_free.restype = None
    L.leon_headers_decode.argtypes = [C.POINTER(HeaderStream), C.c_int32, C.POINTER(C.c_void_p)]
    L.leon_decode_dna.argtypes = [C.POINTER(DnaStream), C.c_int32, C.POINTER(C.c_void_p)]
    L.leon_decode_dna_device.argtypes = [C.POINTER(DnaStream), C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_float)]
    L.leon_decompress_file_device.argtypes = [C.c_char_p, C.c_char_p, C.c_int32, C.c_int32, C.POINTER(FileStats)]
    L.leon_decoded_block.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.c_uint64)]
    L.leon_decoded_free.argtypes = [C.c_void_p]
    L.leon_decoded_free.restype = None
    L.leon_container_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    L.leon_container_close.argtypes = [C.c_void_p]
    L.leon_container_close.restype = None
    L.leon_container_has.argtypes = [C.c_void_p, C.c_char_p]
    L.leon_container_dataset.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.c_uint64)]
    L.leon_container_attr.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.POINTER(C.c_char_p)]
    _lib = L
    return L


def _arr(ptr, n, dtype, copy=True):
    if not n or not ptr:
        return np.zeros(0, dtype)
    a = np.ctypeslib.as_array(ptr, shape=(int(n),))
    return a.astype(dtype, copy=True) if copy else a


def _decoded_lines(L, h, nblocks):
    lines = []
    for b in range(nblocks):
        p, n = C.POINTER(C.c_uint8)(), C.c_uint64()
        L.leon_decoded_block(h, b, C.byref(p), C.byref(n))
        lines += _arr(p, n.value, np.uint8).tobytes().split(b"\n")[:-1]
    return lines


def decode_headers(header, nthreads=0):
    """Inverse of encode_headers: header = dict(first_header, blocksizes, data); returns the names as a list of bytes."""
    L = load_library()
    fh = np.frombuffer(bytes(header["first_header"]), np.uint8)
    bs = np.ascontiguousarray(header["blocksizes"], np.uint64)
    data = np.ascontiguousarray(header["data"], np.uint8)
    hs = HeaderStream()
    hs.first_header = fh.ctypes.data_as(C.POINTER(C.c_uint8)) if len(fh) else None
    hs.first_header_nbytes = len(fh)
    hs.nblocks = len(bs) // 2
    hs.blocksizes = bs.ctypes.data_as(C.POINTER(C.c_uint64))
    hs.data = data.ctypes.data_as(C.POINTER(C.c_uint8))
    hs.data_nbytes = len(data)
    h = C.c_void_p()
    rc = L.leon_headers_decode(C.byref(hs), int(nthreads), C.byref(h))
    if rc != 0:
        raise LeonError(rc, L.leon_last_error(None).decode())
    try:
        return _decoded_lines(L, h, len(bs) // 2)
    finally:
        L.leon_decoded_free(h)


def _dna_stream(result, kmer_size):
    keep = [np.ascontiguousarray(result[k], np.uint8) for k in ("bloom", "anchors_dict", "dna")]
    bs = np.ascontiguousarray(result["blocksizes"], np.uint64)
    # the filter is read 8 bytes at a time: pad it like the device copy
    bloom = np.concatenate([keep[0], np.zeros(16, np.uint8)])
    s = DnaStream(int(kmer_size), int(result.get("max_read_size", -1)), bloom.ctypes.data, len(bloom), int(result["bloom_bits"]),
                  keep[1].ctypes.data, len(keep[1]), int(result["anchor_count"]), len(bs) // 2, bs.ctypes.data,
                  keep[2].ctypes.data, len(keep[2]))
    return s, (keep, bs, bloom)


def decode_dna(result, kmer_size, nthreads=0):
    """Decodes the sequence stream of a result dict (bloom, bloom_bits, anchors_dict, anchor_count, blocksizes, dna,
    max_read_size) back to reads: returns a list of bytes, one per read (C/DnaCoder.cpp:1221-1622)."""
    L = load_library()
    s, keep = _dna_stream(result, kmer_size)
    h = C.c_void_p()
    rc = L.leon_decode_dna(C.byref(s), int(nthreads), C.byref(h))
    if rc != 0:
        raise LeonError(rc, L.leon_last_error(None).decode())
    try:
        return _decoded_lines(L, h, int(s.nblocks))
    finally:
        L.leon_decoded_free(h)


def decode_dna_device(result, kmer_size, device=0, with_time=False):
    """decode_dna on the GPU (k_decode_blocks, one warp per block); result needs max_read_size.  with_time: also the
    kernel's device time in ms."""
    L = load_library()
    s, keep = _dna_stream(result, kmer_size)
    h = C.c_void_p()
    ms = C.c_float(0)
    rc = L.leon_decode_dna_device(C.byref(s), int(device), C.byref(h), C.byref(ms))
    if rc != 0:
        raise LeonError(rc, L.leon_last_error(None).decode())
    try:
        lines = _decoded_lines(L, h, int(s.nblocks))
    finally:
        L.leon_decoded_free(h)
    return (lines, ms.value) if with_time else lines


class Container:
    """Reader of a `.leon` container (ours or one written by the reference through HDF5)."""

    def __init__(self, path):
        self.L = load_library()
        self.h = C.c_void_p()
        rc = self.L.leon_container_open(os.fsencode(path), C.byref(self.h))
        if rc != 0:
            raise LeonError(rc, self.L.leon_last_error(None).decode())

    def close(self):
        if self.h:
            self.L.leon_container_close(self.h)
            self.h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def has(self, path):
        return bool(self.L.leon_container_has(self.h, path.encode()))

    def dataset(self, path):
        p, n = C.POINTER(C.c_uint8)(), C.c_uint64()
        rc = self.L.leon_container_dataset(self.h, path.encode(), C.byref(p), C.byref(n))
        if rc != 0:
            raise LeonError(rc, self.L.leon_last_error(None).decode())
        return _arr(p, n.value, np.uint8)

    def attr(self, dataset_path, name):
        v = C.c_char_p()
        rc = self.L.leon_container_attr(self.h, dataset_path.encode(), name.encode(), C.byref(v))
        if rc != 0:
            raise LeonError(rc, self.L.leon_last_error(None).decode())
        return v.value.decode()


def read_file(path, want_quals=True, want_headers=True):
    """FASTA/FASTQ reader of the file-level path (bank/impl/BankFasta.cpp:484-570): returns dict(bases, offsets, quals,
    headers, header_offsets, is_fastq) as numpy copies."""
    L = load_library()
    fr = C.POINTER(FileReads)()
    rc = L.leon_read_file(os.fsencode(path), int(want_quals), int(want_headers), C.byref(fr))
    if rc != 0:
        raise LeonError(rc, L.leon_last_error(None).decode())
    try:
        f = fr.contents
        n = int(f.reads.nreads)
        offs = np.ctypeslib.as_array(C.cast(f.reads.offsets, C.POINTER(C.c_uint64)), shape=(n + 1,)).copy()
        tot = int(offs[-1])
        u8 = C.POINTER(C.c_uint8)
        out = dict(offsets=offs, bases=_arr(C.cast(f.reads.bases, u8), tot, np.uint8),
                   quals=_arr(C.cast(f.reads.quals, u8), tot, np.uint8) if f.reads.quals else None,
                   is_fastq=bool(f.is_fastq), threads_used=int(f.threads_used), headers=None, header_offsets=None)
        if want_headers:
            ho = np.ctypeslib.as_array(f.header_offsets, shape=(n + 1,)).copy()
            out.update(header_offsets=ho, headers=_arr(f.headers, int(ho[-1]), np.uint8))
        return out
    finally:
        L.leon_file_reads_free(fr)


def compress_file(in_path, out_path=None, kmer_size=31, reads_per_block=50000, lossless=False, abundance=None, noheader=False,
                  noqual=False, seq_only=False, device=0, nthreads=0, stream_reads=0, count_passes=0):
    """`leon -file in_path -c` end to end on the B200 path (C/Leon.cpp:342-470); returns the statistics as a dict."""
    L = load_library()
    o = FileOpts(kmer_size, reads_per_block, int(lossless), int(abundance is not None), int(noheader), int(noqual),
                 int(seq_only), device, nthreads, int(stream_reads), int(count_passes))
    st = FileStats()
    rc = L.leon_compress_file(os.fsencode(in_path), os.fsencode(out_path) if out_path else None, C.byref(o), C.byref(st))
    if rc != 0:
        raise LeonError(rc, L.leon_last_error(None).decode())
    d = {n: getattr(st, n) for n, _ in FileStats._fields_ if not n.startswith("reserved") and n != "gpu"}
    d["gpu"] = {n: getattr(st.gpu, n) for n, _ in Timings._fields_ if n != "reserved"}
    return d


def decompress_file_device(in_path, out_path=None, nthreads=0, device=0):
    """decompress_file with the sequence stream decoded on the GPU (names and qualities on host threads beside it)."""
    L = load_library()
    st = FileStats()
    rc = L.leon_decompress_file_device(os.fsencode(in_path), os.fsencode(out_path) if out_path else None, int(nthreads),
                                       int(device), C.byref(st))
    if rc != 0:
        raise LeonError(rc, L.leon_last_error(None).decode())
    return st


def decompress_file(in_path, out_path=None, nthreads=0):
    """`leon -file in_path -d` (C/Leon.cpp:1311-1470, 1697-1860) on host threads; returns the statistics as a dict."""
    L = load_library()
    st = FileStats()
    rc = L.leon_decompress_file(os.fsencode(in_path), os.fsencode(out_path) if out_path else None, int(nthreads), C.byref(st))
    if rc != 0:
        raise LeonError(rc, L.leon_last_error(None).decode())
    d = {n: getattr(st, n) for n, _ in FileStats._fields_ if not n.startswith("reserved") and n != "gpu"}
    d["dna_ms"], d["qual_ms"] = st.reserved1[0], st.reserved1[1]
    return d


def encode_headers(text, offsets, reads_per_block=50000, nthreads=0):
    """Host header coder (C/HeaderCoder.cpp:257-524 as driven by C/Leon.cpp:786-899): text = the read names without the
    leading '>' / '@', concatenated; offsets = nreads + 1 positions.  Returns dict(first_header, blocksizes, data,
    coded_serially), the `header` argument of write_container."""
    L = load_library()
    text = np.ascontiguousarray(text, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.uint64)
    hs = C.POINTER(HeaderStream)()
    rc = L.leon_headers_encode(text.ctypes.data, offsets.ctypes.data, len(offsets) - 1, int(reads_per_block), int(nthreads), C.byref(hs))
    if rc != 0:
        raise LeonError(rc, L.leon_last_error(None).decode())
    try:
        h = hs.contents
        return dict(first_header=_arr(h.first_header, h.first_header_nbytes, np.uint8).tobytes(),
                    blocksizes=_arr(h.blocksizes, 2 * h.nblocks, np.uint64), data=_arr(h.data, h.data_nbytes, np.uint8),
                    coded_serially=bool(h.coded_serially))
    finally:
        L.leon_headers_free(hs)


def write_container(path, result, kmer_size, is_fasta, header=None):
    """Writes a result of Leon.compress()/fetch() (or any dict with the same keys) as the `.leon` HDF5 container the
    reference's `leon -d` decodes (C/Leon.cpp:485-531, 581-710, 965-1081).  `header` = None for -noheader/-seq-only,
    else a dict(first_header=bytes, blocksizes=uint64[2n], data=uint8[]) from the host header coder."""
    L = load_library()
    keep = []

    def ptr(a, dtype, ctype):
        a = np.ascontiguousarray(a, dtype)
        keep.append(a)
        return a.ctypes.data_as(C.POINTER(ctype))

    r = Result()
    r.bloom, r.bloom_nbytes, r.bloom_bits = ptr(result["bloom"], np.uint8, C.c_uint8), len(result["bloom"]), int(result["bloom_bits"])
    r.anchors_dict, r.anchors_dict_nbytes = ptr(result["anchors_dict"], np.uint8, C.c_uint8), len(result["anchors_dict"])
    r.anchor_count = int(result["anchor_count"])
    r.nblocks = len(result["blocksizes"]) // 2
    r.blocksizes = ptr(result["blocksizes"], np.uint64, C.c_uint64)
    r.dna, r.dna_nbytes = ptr(result["dna"], np.uint8, C.c_uint8), len(result["dna"])
    if not is_fasta:
        if result.get("qual_z") is not None and len(result["qual_z"]):
            r.qual_z, r.qual_z_nbytes = ptr(result["qual_z"], np.uint8, C.c_uint8), len(result["qual_z"])
            r.qual_z_block_sizes = ptr(result["qual_z_block_sizes"], np.uint64, C.c_uint64)
        else:
            r.qual_text, r.qual_nbytes = ptr(result["qual_text"], np.uint8, C.c_uint8), len(result["qual_text"])
            r.qual_block_sizes = ptr(result["qual_block_sizes"], np.uint64, C.c_uint64)
    r.read_count, r.total_dna_size = int(result["read_count"]), int(result["total_dna_size"])
    r.min_read_size, r.max_read_size = int(result["min_read_size"]), int(result["max_read_size"])
    info = ContainerInfo(int(kmer_size), int(bool(is_fasta)), int(header is not None))
    if header is not None:
        fh = np.frombuffer(bytes(header["first_header"]), np.uint8)
        bs = np.ascontiguousarray(header["blocksizes"], np.uint64)
        hd = np.ascontiguousarray(header["data"], np.uint8)
        keep += [fh, bs, hd]
        info.first_header, info.first_header_nbytes = fh.ctypes.data, len(fh)
        info.header_nblocks, info.header_blocksizes, info.header_data = len(bs) // 2, bs.ctypes.data, hd.ctypes.data
    rc = L.leon_write_container(os.fsencode(path), C.byref(r), C.byref(info))
    if rc != 0:
        raise LeonError(rc, L.leon_last_error(None).decode())


class PinnedArray:
    """numpy view over pinned host memory obtained from leon_alloc_pinned."""

    def __init__(self, nbytes, dtype=np.uint8):
        L = load_library()
        self._L = L
        self.nbytes = int(nbytes)
        self.ptr = L.leon_alloc_pinned(max(self.nbytes, 1))
        if not self.ptr:
            raise LeonError(-3, "leon_alloc_pinned(%d) failed" % nbytes)
        buf = (C.c_uint8 * max(self.nbytes, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=np.uint8, count=self.nbytes).view(dtype)

    def free(self):
        if self.ptr:
            self.array = None
            self._L.leon_free_pinned(self.ptr)
            self.ptr = None


class Leon:
    """Host-side mirror of the reference's `Leon` tool for the sequence stream of `leon -c`.

    Option names follow the command line of the reference (C/Leon.cpp:102-140): kmer_size = -kmer-size,
    reads_per_block = -reads, lossless = -lossless, abundance = -abundance (only its presence matters, as in the
    reference), seq_only/noqual decide whether qualities are kept.
    """

    def __init__(self, kmer_size=31, reads_per_block=50000, lossless=False, abundance=None, has_qual=False, device=0,
                 anchor_batch=0, dict_on_device=False, qual_zlib_on_device=False, skip_qual_text=False, skip_read_arrays=False):
        self.L = load_library()
        self.opts = Opts(kmer_size, reads_per_block, int(has_qual), int(lossless), int(abundance is not None), device,
                         anchor_batch, int(dict_on_device), int(qual_zlib_on_device), int(skip_qual_text), int(skip_read_arrays))
        self.ctx = C.c_void_p()
        rc = self.L.leon_create(C.byref(self.opts), C.byref(self.ctx))
        if rc != 0:
            raise LeonError(rc, self.L.leon_last_error(None).decode())
        self._keep = None

    def close(self):
        if self.ctx:
            self.L.leon_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise LeonError(rc, self.L.leon_last_error(self.ctx).decode())

    def _reads(self, bases, offsets, quals):
        bases = np.ascontiguousarray(bases, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        q = np.ascontiguousarray(quals, np.uint8) if (self.opts.has_qual and quals is not None) else None
        self._keep = (bases, offsets, q)
        return Reads(bases.ctypes.data, offsets.ctypes.data, q.ctypes.data if q is not None else None, len(offsets) - 1)

    @staticmethod
    def _result(r, copy=True):
        hq = r.qual_nbytes > 0 or r.qual_z_nbytes > 0
        return dict(
            qual_z=_arr(r.qual_z, r.qual_z_nbytes, np.uint8, copy),
            qual_z_block_sizes=_arr(r.qual_z_block_sizes, r.nblocks if r.qual_z_nbytes else 0, np.uint64, copy),
            cutoff=int(r.count.cutoff), nb_solids_cutoff=int(r.count.nb_solids_cutoff),
            nb_solid_total=int(r.count.nb_solid_total), nb_distinct=int(r.count.nb_distinct), nb_kmers=int(r.count.nb_kmers),
            histogram=_arr(r.count.histogram, 10001, np.uint64, copy),
            bloom=_arr(r.bloom, r.bloom_nbytes, np.uint8, copy), bloom_bits=int(r.bloom_bits),
            anchors_dict=_arr(r.anchors_dict, r.anchors_dict_nbytes, np.uint8, copy), anchor_count=int(r.anchor_count),
            nblocks=int(r.nblocks), blocksizes=_arr(r.blocksizes, 2 * r.nblocks, np.uint64, copy),
            dna=_arr(r.dna, r.dna_nbytes, np.uint8, copy),
            qual_text=_arr(r.qual_text, r.qual_nbytes, np.uint8, copy),
            qual_block_sizes=_arr(r.qual_block_sizes, r.nblocks if hq else 0, np.uint64, copy),
            read_count=int(r.read_count), total_dna_size=int(r.total_dna_size),
            min_read_size=int(r.min_read_size), max_read_size=int(r.max_read_size),
            anchor_pos=_arr(r.anchor_pos, r.read_count if r.anchor_pos else 0, np.int32, copy),
            anchor_addr=_arr(r.anchor_addr, r.read_count if r.anchor_addr else 0, np.uint32, copy),
            reads_without_anchor=int(r.reads_without_anchor),
        )

    # one-shot, host buffers in / host buffers out
    def compress(self, bases, offsets, quals=None, copy=True):
        rd = self._reads(bases, offsets, quals)
        r = Result()
        self._check(self.L.leon_compress(self.ctx, C.byref(rd), C.byref(r)))
        return self._result(r, copy)

    # phase by phase
    def upload(self, bases, offsets, quals=None):
        rd = self._reads(bases, offsets, quals)
        self._check(self.L.leon_upload(self.ctx, C.byref(rd)))

    def count(self):
        r = CountResult()
        self._check(self.L.leon_count_resident(self.ctx, C.byref(r)))
        return dict(cutoff=int(r.cutoff), nb_solids_cutoff=int(r.nb_solids_cutoff), nb_solid_total=int(r.nb_solid_total),
                    nb_distinct=int(r.nb_distinct), nb_kmers=int(r.nb_kmers), histogram=_arr(r.histogram, 10001, np.uint64))

    def bloom_build(self):
        bits = C.POINTER(C.c_uint8)()
        nbytes, red = C.c_uint64(), C.c_uint64()
        self._check(self.L.leon_bloom_build(self.ctx, C.byref(bits), C.byref(nbytes), C.byref(red)))
        return _arr(bits, nbytes.value, np.uint8), int(red.value)

    def encode(self):
        r = Result()
        self._check(self.L.leon_encode(self.ctx, C.byref(r)))
        return self._result(r)

    # streaming path (include/leon_b200.h "Streaming path"): batches of reads, counted in passes, then encoded in file order
    def count_begin(self, total_reads, total_bases, count_passes=0):
        plan = StreamPlan(int(total_reads), int(total_bases), int(count_passes))
        npass = C.c_uint32()
        self._check(self.L.leon_count_begin(self.ctx, C.byref(plan), C.byref(npass)))
        return int(npass.value)

    def _batch(self, bases, offsets, quals):
        """A Reads struct over the caller's arrays WITHOUT copies (they must be contiguous u8 / u64 and stay alive)."""
        assert bases.dtype == np.uint8 and offsets.dtype == np.uint64 and bases.flags.c_contiguous and offsets.flags.c_contiguous
        q = quals if (self.opts.has_qual and quals is not None) else None
        return Reads(bases.ctypes.data, offsets.ctypes.data, q.ctypes.data if q is not None else None, len(offsets) - 1), (bases, offsets, q)

    def prefetch(self, bases, offsets, quals=None):
        """Starts the upload of the batch that will be submitted next (leon_stream_prefetch); the arrays must be the very
        ones passed to the next count_batch / encode_block call."""
        rd, keep = self._batch(bases, offsets, quals)
        self._pf_keep = (getattr(self, "_pf_keep", ()) + (keep,))[-3:]      # two announcements may be open + the batch in work
        self._check(self.L.leon_stream_prefetch(self.ctx, C.byref(rd), int(quals is not None)))

    def count_batch(self, bases, offsets, last):
        rd, keep = self._batch(np.ascontiguousarray(bases, np.uint8), np.ascontiguousarray(offsets, np.uint64), None)
        self._check(self.L.leon_count(self.ctx, C.byref(rd), int(bool(last))))

    def count_finish(self):
        r = CountResult()
        self._check(self.L.leon_count_finish(self.ctx, C.byref(r)))
        return dict(cutoff=int(r.cutoff), nb_solids_cutoff=int(r.nb_solids_cutoff), nb_solid_total=int(r.nb_solid_total),
                    nb_distinct=int(r.nb_distinct), nb_kmers=int(r.nb_kmers), histogram=_arr(r.histogram, 10001, np.uint64))

    def encode_block(self, first_block, bases, offsets, quals=None, copy=True):
        """Codes the blocks of one batch; returns their outputs (views into the context's buffers when copy=False: valid until
        the next call)."""
        rd, keep = self._batch(np.ascontiguousarray(bases, np.uint8), np.ascontiguousarray(offsets, np.uint64),
                               np.ascontiguousarray(quals, np.uint8) if quals is not None else None)
        bo = BlockOut()
        self._check(self.L.leon_encode_block(self.ctx, int(first_block), C.byref(rd), C.byref(bo)))
        nb = int(bo.nblocks)
        hq = bo.qual_nbytes > 0 or bo.qual_z_nbytes > 0
        return dict(first_block=int(bo.first_block), nblocks=nb, read_count=int(bo.read_count),
                    dna=_arr(bo.dna, bo.dna_nbytes, np.uint8, copy), blocksizes=_arr(bo.blocksizes, 2 * nb, np.uint64, copy),
                    qual_text=_arr(bo.qual_text, bo.qual_nbytes, np.uint8, copy),
                    qual_block_sizes=_arr(bo.qual_block_sizes, nb if hq else 0, np.uint64, copy),
                    qual_z=_arr(bo.qual_z, bo.qual_z_nbytes, np.uint8, copy),
                    qual_z_block_sizes=_arr(bo.qual_z_block_sizes, nb if bo.qual_z_nbytes else 0, np.uint64, copy),
                    anchor_pos=_arr(bo.anchor_pos, bo.read_count if bo.anchor_pos else 0, np.int32, copy),
                    anchor_addr=_arr(bo.anchor_addr, bo.read_count if bo.anchor_addr else 0, np.uint32, copy))

    def encode_finish(self, copy=True):
        r = Result()
        self._check(self.L.leon_encode_finish(self.ctx, C.byref(r)))
        return self._result(r, copy)

    def compress_stream(self, bases, offsets, quals=None, batch_reads=None, count_passes=0, batches=None, prefetch="alternate"):
        """The whole set through the streaming entry points, `batch_reads` reads at a time (or the explicit list of (lo, hi) read
        ranges `batches`); returns the result dict of compress() assembled from the per-batch outputs, plus `count_passes`.
        `prefetch`: "all" announces every next batch before the current one is submitted (what a caller after speed does),
        "alternate" every other one in the encoding pass (both upload paths of the library stay exercised), "none" never."""
        offsets = np.ascontiguousarray(offsets, np.uint64)
        n = len(offsets) - 1
        if batches is None:
            batches = [(lo, min(n, lo + batch_reads)) for lo in range(0, n, batch_reads)] or [(0, 0)]

        bases = np.ascontiguousarray(bases, np.uint8)
        quals = np.ascontiguousarray(quals, np.uint8) if (quals is not None and self.opts.has_qual) else None
        cuts = []
        for lo, hi in batches:          # the arrays of every batch once, so that a prefetched batch is recognised by its pointers
            b0, b1 = int(offsets[lo]), int(offsets[hi])
            cuts.append((bases[b0:b1], np.ascontiguousarray(offsets[lo:hi + 1] - offsets[lo]), quals[b0:b1] if quals is not None else None))

        npass = self.count_begin(n, int(offsets[-1]), count_passes)
        for ps in range(npass):
            for i, (b, o, _q) in enumerate(cuts):
                if i + 1 < len(cuts) and prefetch != "none":
                    self.prefetch(cuts[i + 1][0], cuts[i + 1][1])
                elif prefetch == "all":     # behind the last batch of a pass: the first one of the next pass / of the encoding pass
                    if ps + 1 < npass:
                        self.prefetch(cuts[0][0], cuts[0][1])
                    else:
                        self.prefetch(*cuts[0])
                self.count_batch(b, o, i == len(cuts) - 1)
        self.count_finish()
        self.bloom_build()
        keys = ("dna", "blocksizes", "qual_text", "qual_block_sizes", "qual_z", "qual_z_block_sizes", "anchor_pos", "anchor_addr")
        parts = {k: [] for k in keys}
        first = 0
        for i, (b, o, q) in enumerate(cuts):
            if i + 1 < len(cuts) and (prefetch == "all" or (prefetch == "alternate" and i % 2 == 0)):
                self.prefetch(*cuts[i + 1])
            bo = self.encode_block(first, b, o, q)
            for k in keys:
                parts[k].append(bo[k])
            first += bo["nblocks"]
        out = self.encode_finish()
        for k, v in parts.items():
            out[k] = np.concatenate(v) if v else np.zeros(0, np.uint8)
        out["nblocks"] = first
        out["count_passes"] = npass
        return out

    def run_device(self):
        self._check(self.L.leon_run_device(self.ctx))

    def fetch(self, copy=True):
        r = Result()
        self._check(self.L.leon_fetch(self.ctx, C.byref(r)))
        return self._result(r, copy)

    def timings(self):
        t = Timings()
        self._check(self.L.leon_get_timings(self.ctx, C.byref(t)))
        return {n: getattr(t, n) for n, _ in Timings._fields_ if n != "reserved"}

    def kernel_times(self):
        """{kernel name: (total ms, launches)} of the last run."""
        arr = (KernelTime * 64)()
        n = C.c_int()
        self._check(self.L.leon_get_kernel_times(self.ctx, arr, 64, C.byref(n)))
        return {arr[i].name.decode(): (float(arr[i].ms_total), int(arr[i].launches)) for i in range(min(n.value, 64))}

    @staticmethod
    def comm_unique_id():
        """128 opaque bytes (an ncclUniqueId) created by rank 0 and handed to every rank."""
        L = load_library()
        buf = (C.c_uint8 * 128)()
        rc = L.leon_comm_unique_id(buf)
        if rc != 0:
            raise LeonError(rc, L.leon_last_error(None).decode())
        return bytes(buf)

    def comm_init(self, rank, nranks, unique_id):
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        self._check(self.L.leon_comm_init(self.ctx, rank, nranks, buf))
        self.rank, self.nranks = rank, nranks

    def comm_batch_reads(self):
        v = C.c_uint64()
        self._check(self.L.leon_comm_batch_reads(self.ctx, C.byref(v)))
        return int(v.value)

    def set_stream(self, cuda_stream_ptr):
        self._check(self.L.leon_set_stream(self.ctx, C.c_void_p(cuda_stream_ptr)))

    def debug_bloom_query(self, kmers_lo_hi, right):
        k = np.ascontiguousarray(kmers_lo_hi, np.uint64).reshape(-1, 2)
        n = len(k)
        o4, o1 = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
        self._check(self.L.leon_debug_bloom_query(self.ctx, k.ctypes.data, n, int(right), o4.ctypes.data, o1.ctypes.data))
        return o4, o1

    def debug_hashes(self, kmers_lo_hi):
        k = np.ascontiguousarray(kmers_lo_hi, np.uint64).reshape(-1, 2)
        out = np.zeros((len(k), 12), np.uint64)
        self._check(self.L.leon_debug_hashes(self.ctx, k.ctypes.data, len(k), out.ctypes.data))
        return out

    def debug_fetch(self, name, dtype=np.uint8):
        nb = C.c_uint64()
        self._check(self.L.leon_debug_fetch(self.ctx, name.encode(), None, 0, C.byref(nb)))
        out = np.zeros(nb.value, np.uint8)
        if nb.value:
            self._check(self.L.leon_debug_fetch(self.ctx, name.encode(), out.ctypes.data, nb.value, C.byref(nb)))
        return out.view(dtype)
