"""Host-side sanitizer tier: the I/O tests, including the corrupted-input fuzz, against an
AddressSanitizer + UBSan build of the C++ I/O library, then the I/O tests against a ThreadSanitizer build (the
shuffling prefetcher and the record queue are threaded) -- scripts/asan_io.sh."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(600)
def test_io_library_is_clean_under_asan_ubsan_and_tsan():
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    asan = subprocess.run(["gcc", "-print-file-name=libasan.so"], capture_output=True, text=True).stdout.strip()
    if not os.path.isabs(asan) or not os.path.exists(asan):
        pytest.skip("libasan is not installed")
    if os.environ.get("DTF_IO_LIB"):
        pytest.skip("already running against another build of the library")
    env = {k: v for k, v in os.environ.items() if k not in ("LD_PRELOAD", "ASAN_OPTIONS")}
    p = subprocess.run(["bash", os.path.join(ROOT, "scripts", "asan_io.sh")], capture_output=True, text=True,
                       env=env, timeout=580)
    tail = (p.stdout + p.stderr)[-3000:]
    assert p.returncode == 0 and p.stdout.count(" passed") == 2, tail  # both stages ran their tests
    assert "ERROR: AddressSanitizer" not in tail and "runtime error:" not in tail, tail
    assert "WARNING: ThreadSanitizer" not in tail, tail
