"""leon_b200 — B200-native Leon compression hot path.

The product is libleon_b200.so (hand-written sm_100a CUDA behind the C ABI of include/leon_b200.h).
This package is the thin Python host mirror used by the parity tests and bench.py: it binds the C ABI with
ctypes and keeps the names of the reference's interface (`-kmer-size`, `-reads`, `-lossless`, `-abundance`, ...).
It contains no compute and no fallback: if the library or a CUDA device is missing, calls raise.
"""
import os as _os

# The path runs its stages on about 15 streams; with the default 8 hardware connections unrelated streams share a queue and a
# 38 ms range coder kernel holds up whatever is queued behind it (measured on B200: 111 vs 82 ms per configs[2] run).  Costs
# 0.5-1 s of context creation, so the one-shot CLI does not set it.  Only effective before the CUDA context is created.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

from .api import Leon, LeonError, load_library, write_container, encode_headers, read_file, compress_file, decompress_file, decompress_file_device, Container, decode_dna, decode_dna_device, decode_headers, ABI_SYMBOLS  # noqa: F401
