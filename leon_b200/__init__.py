"""leon_b200 — B200-native Leon compression hot path.

The product is libleon_b200.so (hand-written sm_100a CUDA behind the C ABI of include/leon_b200.h).
This package is the thin Python host mirror used by the parity tests and bench.py: it binds the C ABI with
ctypes and keeps the names of the reference's interface (`-kmer-size`, `-reads`, `-lossless`, `-abundance`, ...).
It contains no compute and no fallback: if the library or a CUDA device is missing, calls raise.
"""
import os as _os

# The path runs its stages on a dozen streams; with the default 8 hardware connections unrelated streams share a queue.
# Only effective when set before the CUDA context of the process is created.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

from .api import Leon, LeonError, load_library, write_container, encode_headers, read_file, compress_file, decompress_file, decompress_file_device, Container, decode_dna, decode_dna_device, decode_headers, ABI_SYMBOLS  # noqa: F401
