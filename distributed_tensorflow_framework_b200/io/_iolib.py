"""ctypes binding of libdtf_b200_io.so (CPU-side native I/O; see io/csrc/dtf_io.cc)."""
import ctypes
import os

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# DTF_IO_LIB: load another build of the library instead (the ASan / UBSan build of scripts/asan_io.sh)
_LIB_PATH = os.environ.get("DTF_IO_LIB") or os.path.join(_PKG, "lib", "libdtf_b200_io.so")
_lib = None

c_void_p, c_char_p, c_int, c_ull, c_ll = ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_ulonglong, ctypes.c_longlong
P = ctypes.POINTER

_SIGS = {
    "dtf_io_version": (c_int, []),
    "dtf_crc32c": (ctypes.c_uint, [c_void_p, c_ull]),
    "dtf_crc32c_masked": (ctypes.c_uint, [c_void_p, c_ull]),
    "dtf_tfrecord_open": (c_void_p, [c_char_p, c_int]),
    "dtf_tfrecord_next": (c_int, [c_void_p, P(c_void_p), P(c_ull)]),
    "dtf_tfrecord_close": (None, [c_void_p]),
    "dtf_tfrecord_writer_open": (c_void_p, [c_char_p]),
    "dtf_tfrecord_write": (c_int, [c_void_p, c_void_p, c_ull]),
    "dtf_tfrecord_writer_close": (c_int, [c_void_p]),
    "dtf_example_count": (c_int, [c_void_p, c_ull, c_char_p, P(c_int)]),
    "dtf_example_bytes": (c_int, [c_void_p, c_ull, c_char_p, c_int, P(c_ull), P(c_ull)]),
    "dtf_example_floats": (c_int, [c_void_p, c_ull, c_char_p, c_void_p, c_int]),
    "dtf_example_int64s": (c_int, [c_void_p, c_ull, c_char_p, c_void_p, c_int]),
    "dtf_prefetch_open": (c_void_p, [c_char_p, c_int, c_int, c_int, c_ull, c_ull]),
    "dtf_prefetch_next": (c_int, [c_void_p, P(c_void_p), P(c_ull)]),
    "dtf_prefetch_close": (None, [c_void_p]),
    "dtf_bundle_writer_open": (c_void_p, [c_char_p, c_int]),
    "dtf_bundle_writer_add": (c_int, [c_void_p, c_char_p, c_int, P(c_ll), c_int, c_void_p, c_ull, c_int]),
    "dtf_bundle_writer_finish": (c_int, [c_void_p]),
    "dtf_bundle_reader_open": (c_void_p, [c_char_p]),
    "dtf_bundle_reader_num_tensors": (c_int, [c_void_p]),
    "dtf_bundle_reader_num_shards": (c_int, [c_void_p]),
    "dtf_bundle_reader_entry": (c_int, [c_void_p, c_int, c_char_p, c_int, P(c_int), P(c_ll), P(c_int), P(c_ll), P(c_int)]),
    "dtf_bundle_reader_read": (c_int, [c_void_p, c_char_p, c_void_p, c_ull]),
    "dtf_bundle_reader_close": (None, [c_void_p]),
    "dtf_events_open": (c_void_p, [c_char_p, ctypes.c_double]),
    "dtf_events_scalar": (c_int, [c_void_p, c_char_p, ctypes.c_float, c_ll, ctypes.c_double]),
    "dtf_events_flush": (c_int, [c_void_p]),
    "dtf_events_close": (c_int, [c_void_p]),
}


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            from .. import build
            build.build_io()
        _lib = ctypes.CDLL(_LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(_lib, name)
            fn.restype = res
            fn.argtypes = args
    return _lib
