#!/bin/bash
# tools/build_ref_apps.sh -- build the reference's llama-tts-mio and mio-tts-server on top of libmio_tts_b200.so.
# Needs /root/reference (this container); output: tools/ref_apps/_build/bin/{llama-tts-mio,mio-tts-server,lib*.so}
# (git-ignored, travels to the GPU box with the snapshot).  tests/test_ref_apps.py and tools/server_bench.py use it.
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
REF="${REF:-/root/reference}"
LIB="$ROOT/mio_tts_cpp_b200/libmio_tts_b200.so"
[ -f "$LIB" ] || make -C "$ROOT/mio_tts_cpp_b200/csrc" -j8
B="$ROOT/tools/ref_apps/_build"
mkdir -p "$B"
# the copy next to the binaries is what $ORIGIN resolves; refreshed on every build
cmake -S "$ROOT/tools/ref_apps" -B "$B" -DCMAKE_BUILD_TYPE=Release -DREF="$REF" -DMIO_B200_LIB="$LIB" > "$B/configure.log" 2>&1 \
  || { tail -30 "$B/configure.log"; exit 1; }
cmake --build "$B" -j"${JOBS:-8}" --target llama-tts-mio mio-tts-server llama-tts-mio-ref > "$B/build.log" 2>&1 || { tail -40 "$B/build.log"; exit 1; }
# real files under the sonames the binaries ask for (symlinks may not survive the snapshot to the GPU box);
# libmio_tts_b200.so itself is found through LD_LIBRARY_PATH=<repo>/mio_tts_cpp_b200 (tests/test_ref_apps.py)
for f in "$B"/bin/lib*.so.0; do cp --remove-destination "$(readlink -f "$f")" "$f"; done
find "$B/bin" \( -name 'lib*.so' -o -name 'lib*.so.0.*' \) -delete
# keep only what has to travel
find "$B" -mindepth 1 -maxdepth 1 ! -name bin ! -name '*.log' -exec rm -rf {} +
ls -la "$B/bin"
