#!/bin/bash
# Host-side sanitizer tier: the C++ I/O library (crc32c, TFRecord, Example parser, shuffling prefetcher,
# TensorBundle reader / writer, event files) built with AddressSanitizer + UBSan, then the I/O tests -- including the
# corrupted-input fuzz -- run against that build. Leak detection is off (CPython itself "leaks" at exit).
set -e
cd "$(dirname "$0")/.."
LIB=$(python -c "from distributed_tensorflow_framework_b200 import build; print(build.build_io_sanitized())")
export DTF_IO_LIB="$LIB"
export LD_PRELOAD="$(gcc -print-file-name=libasan.so):$(gcc -print-file-name=libubsan.so)"
export ASAN_OPTIONS="detect_leaks=0:abort_on_error=1:halt_on_error=1"
export UBSAN_OPTIONS="print_stacktrace=1:halt_on_error=1"
python -m pytest tests/test_io.py tests/test_io_fuzz.py -x -q -p no:cacheprovider "$@"
# second stage: ThreadSanitizer build for the threaded pieces (shuffling prefetcher, record queue)
TSAN_LIB=distributed_tensorflow_framework_b200/lib/libdtf_b200_io_tsan.so
g++ -O1 -g -std=c++17 -fPIC -shared -pthread -fsanitize=thread -o "$TSAN_LIB" distributed_tensorflow_framework_b200/io/csrc/*.cc
export DTF_IO_LIB="$PWD/$TSAN_LIB"
export LD_PRELOAD="$(gcc -print-file-name=libtsan.so)"
export TSAN_OPTIONS="halt_on_error=1:report_signal_unsafe=0:exitcode=66"
unset ASAN_OPTIONS UBSAN_OPTIONS
python -m pytest tests/test_io.py -x -q -p no:cacheprovider "$@"
