#!/usr/bin/env bash
# TEST INFRASTRUCTURE.  Builds the UNMODIFIED reference (/root/reference = GATB/leon + gatb-core submodule) and
# drops the two binaries the harness uses into oracle/_ref/ (git-ignored, travels to the GPU box):
#   oracle/_ref/leon          the reference compressor (CPU baseline of bench.py --impl reference; golden fixtures)
#   oracle/_ref/leon64        same, compiled with KSIZE_LIST="64 96 128 160" (k up to 63)
#   oracle/_ref/gatb-h5dump   dataset extractor used by tests/golden/make_golden.py
# The reference does not compile from "a few source files": it needs its vendored HDF5 (configure-generated
# headers) and gatb-core's generated config.hpp, so its own CMake project is driven here, out of tree, in
# /tmp.  Nothing is copied from the reference SOURCES into this repository.  Only runs where /root/reference exists.
set -euo pipefail
REF=${REF:-/root/reference}
OUT="$(cd "$(dirname "$0")" && pwd)/_ref"
[ -d "$REF" ] || { echo "no $REF here: keeping whatever is in $OUT"; exit 0; }
mkdir -p "$OUT"
build() { # dir, extra cmake args...
  local d=$1; shift
  if [ ! -x "$d/bin/leon" ]; then
    mkdir -p "$d" && (cd "$d" && cmake -DCMAKE_POLICY_VERSION_MINIMUM=3.5 -DCMAKE_BUILD_TYPE=Release "$@" "$REF" >cmake.log 2>&1 \
      && make -j"$(nproc)" >make.log 2>&1)
  fi
}
build /tmp/leon_build
cp -f /tmp/leon_build/bin/leon "$OUT/leon"
cp -f /tmp/leon_build/ext/gatb-core/bin/Release/gatb-h5dump "$OUT/gatb-h5dump"
if [ "${WITH_K64:-1}" = 1 ]; then
  build /tmp/leon_build64 -DKSIZE_LIST="64 96 128 160"
  cp -f /tmp/leon_build64/bin/leon "$OUT/leon64"
fi
ls -la "$OUT"
