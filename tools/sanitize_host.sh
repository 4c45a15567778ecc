#!/bin/bash
# Host-side sanitizer passes over the C++ input pipeline (the GPU pool refuses compute-sanitizer; this part needs no GPU):
#   tools/sanitize_host.sh [out_dir]     -> <out_dir>/sanitize_host_{tsan,asan_ubsan,fuzz}_r2.log
# Builds tools/loader_stress.cpp together with csrc/io/{record_loader,crc32c}.cpp under -fsanitize=thread and under
# -fsanitize=address,undefined, writes an MNIST-shaped TFRecord file and runs the stress driver on it.
set -u
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
OUT="${1:-$ROOT/gpurun_out}"
TMP="$(mktemp -d)"
mkdir -p "$OUT"
PYTHONPATH="$ROOT" python - "$TMP/d.tfrecords" <<'PY'
import sys
import numpy as np
from distributed_tensorflow_ibm_mnist_b200.formats import tfrecord
rng = np.random.default_rng(0)
tfrecord.write_mnist_tfrecord(sys.argv[1], [rng.integers(0, 256, 784, dtype=np.uint8).tobytes() for _ in range(6000)],
                              [int(i % 10) for i in range(6000)])
raw = bytearray(open(sys.argv[1], "rb").read())
raw[len(raw) // 2] ^= 0x5A
open(sys.argv[1] + ".corrupt", "wb").write(bytes(raw))
PY
SRCS="$ROOT/tools/loader_stress.cpp $ROOT/csrc/io/record_loader.cpp $ROOT/csrc/io/crc32c.cpp"
INC="-I/usr/local/cuda/include"
rc=0
for mode in tsan asan_ubsan; do
  if [ $mode = tsan ]; then FLAGS="-fsanitize=thread"; else FLAGS="-fsanitize=address,undefined -fno-sanitize-recover=undefined"; fi
  LOG="$OUT/sanitize_host_${mode}_r2.log"
  {
    echo "== g++ -std=c++17 -O1 -g $FLAGS loader_stress.cpp record_loader.cpp crc32c.cpp"
    if g++ -std=c++17 -O1 -g $FLAGS $INC -o "$TMP/stress_$mode" $SRCS -lpthread 2>&1 | grep -v "warning:" ; then :; fi
    if [ -x "$TMP/stress_$mode" ]; then
      TSAN_OPTIONS="halt_on_error=0 exitcode=66" ASAN_OPTIONS="detect_leaks=1 exitcode=67" "$TMP/stress_$mode" "$TMP/d.tfrecords" "$TMP/d.tfrecords.corrupt" 2>&1
      code=$?
      echo "== exit code $code ($([ $code -eq 0 ] && echo 'no reports' || echo 'REPORTS ABOVE'))"
      [ $code -ne 0 ] && rc=1
    else
      echo "== build failed"; rc=2
    fi
  } > "$LOG" 2>&1
  tail -3 "$LOG"
done
# third pass: parser fuzz under ASan + UBSan (every input in an exact-size heap buffer)
LOG="$OUT/sanitize_host_fuzz_r2.log"
{
  echo "== g++ -std=c++17 -O1 -g -fsanitize=address,undefined parse_fuzz.cpp crc32c.cpp"
  g++ -std=c++17 -O1 -g -fsanitize=address,undefined -fno-sanitize-recover=undefined $INC -o "$TMP/fuzz" "$ROOT/tools/parse_fuzz.cpp" \
      "$ROOT/csrc/io/crc32c.cpp" -lpthread 2>&1 | grep -v "warning:"
  if [ -x "$TMP/fuzz" ]; then
    ASAN_OPTIONS="detect_leaks=1 exitcode=67" "$TMP/fuzz" 300000 2>&1
    code=$?
    echo "== exit code $code ($([ $code -eq 0 ] && echo 'no reports' || echo 'REPORTS ABOVE'))"
    [ $code -ne 0 ] && rc=1
  else
    echo "== build failed"; rc=2
  fi
} > "$LOG" 2>&1
tail -2 "$LOG"
rm -rf "$TMP"
exit $rc
