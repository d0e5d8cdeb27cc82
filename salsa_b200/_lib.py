"""ctypes binding of libsalsa_b200.so (the C ABI declared in include/salsa_b200.h).

There is no CPU fallback: if the shared library is missing, or no CUDA device
is present, every compute call raises.  The library is built in-tree by
``__graft_entry__.build()`` / ``make -C salsa_b200/csrc``.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('SALSA_B200_LIB') or os.path.join(_HERE, 'libsalsa_b200.so')   # override: A/B builds
_lib = None

c_void_p, c_size_t, c_u64, c_i32, c_i64 = C.c_void_p, C.c_size_t, C.c_uint64, C.c_int32, C.c_int64


class IngestStats(C.Structure):
    _fields_ = [('err_offset', C.c_uint64), ('err_code', C.c_int32), ('lines', C.c_uint64), ('pairs', C.c_uint64)]


class SalsaError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__('libsalsa_b200 error %d: %s' % (code, msg))
        self.code = code


DATA_TABLE_FULL = 6
DATA_ERRORS = {1: 'IndexError', 2: 'ValueError', 3: 'KeyError', 4: 'ZeroDivisionError', 5: 'LineTooLong', 6: 'TableFull'}


class DataError(RuntimeError):
    """The BED made the reference's Python raise (or hit a library limit)."""

    def __init__(self, code, offset):
        super().__init__('%s at BED byte offset %d' % (DATA_ERRORS.get(code, 'error %d' % code), offset))
        self.code = code
        self.offset = offset


# name -> (restype, argtypes); every symbol include/salsa_b200.h declares
SIGNATURES = {
    'salsa_last_error': (C.c_char_p, []),
    'salsa_version': (C.c_int, []),
    'salsa_ctx_create': (C.c_int, [C.c_int, C.POINTER(c_void_p)]),
    'salsa_ctx_destroy': (None, [c_void_p]),
    'salsa_ctx_scan_mode': (C.c_int, [c_void_p, C.c_int]),
    'salsa_timers_enable': (C.c_int, [c_void_p, C.c_int]),
    'salsa_timers': (C.c_int, [c_void_p, c_void_p, c_void_p, c_i32, C.c_int]),
    'salsa_names_build': (C.c_int, [c_void_p, C.c_char_p, c_void_p, c_i32, c_void_p, c_void_p]),
    'salsa_names_count': (c_i32, [c_void_p]),
    'salsa_avoid_set': (C.c_int, [c_void_p, c_void_p, c_i32]),
    'salsa_ingest_stats_reset': (C.c_int, [c_void_p, c_void_p]),
    'salsa_ingest_stats_read': (C.c_int, [c_void_p, C.POINTER(IngestStats), c_void_p]),
    'salsa_bed_ingest_dense': (C.c_int, [c_void_p, c_void_p, c_size_t, c_u64, c_void_p, c_void_p, c_void_p]),
    'salsa_sparse_reset': (C.c_int, [c_void_p, c_void_p, c_size_t, c_void_p]),
    'salsa_bed_ingest_sparse': (C.c_int, [c_void_p, c_void_p, c_size_t, c_u64, c_void_p, c_void_p, c_size_t, c_void_p, c_void_p]),
    'salsa_bed_ingest_dense_host': (C.c_int, [c_void_p, c_void_p, c_size_t, c_void_p, c_size_t, c_void_p, c_void_p, c_void_p]),
    'salsa_links_compact_scratch_bytes': (c_size_t, [c_size_t]),
    'salsa_links_compact': (C.c_int, [c_void_p, c_void_p, c_size_t, c_void_p, c_void_p, c_size_t, c_void_p, C.POINTER(c_u64), c_void_p]),
    'salsa_sort_pairs_u64': (C.c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p, C.POINTER(c_size_t), c_void_p]),
    'salsa_links_partition': (C.c_int, [c_void_p, c_void_p, c_size_t, c_i32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    'salsa_sparse_add': (C.c_int, [c_void_p, c_void_p, c_size_t, c_void_p, c_void_p, c_size_t, c_void_p, c_void_p]),
    'salsa_sparse_segment_bytes': (c_size_t, [c_size_t]),
    'salsa_sparse_partition': (C.c_int, [c_void_p, c_void_p, c_size_t, c_i32, c_size_t, c_void_p, c_void_p, c_void_p]),
    'salsa_sparse_add_segments': (C.c_int, [c_void_p, c_i32, c_size_t, c_void_p, c_void_p, c_size_t, c_void_p, c_void_p]),
    'salsa_clean_bed_plan': (C.c_int, [c_void_p, c_void_p, c_size_t, c_void_p, c_void_p, c_size_t, c_void_p, c_void_p, c_void_p]),
    'salsa_clean_bed_emit': (C.c_int, [c_void_p, c_void_p, c_size_t, c_void_p, c_void_p, c_size_t, c_void_p, c_void_p, c_void_p, c_void_p]),
    'salsa_scaffold_re': (C.c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_i32, c_void_p, c_void_p, c_void_p]),
    'salsa_g12': (C.c_int, [c_void_p, c_size_t, c_void_p, c_void_p, c_void_p, c_void_p]),
    'salsa_links_score': (C.c_int, [c_void_p, c_void_p, c_size_t, c_i32] + [c_void_p] * 10),
    'salsa_scaled_scratch_bytes': (c_size_t, [c_i32]),
    'salsa_scaled_scores': (C.c_int, [c_void_p, c_void_p, c_void_p, c_size_t, c_i32] + [c_void_p] * 9),
    'salsa_scaled_scores_static_max_nodes': (c_i32, []),
    'salsa_scaled_scores_static': (C.c_int, [c_void_p, c_void_p, c_void_p, c_size_t, c_i32] + [c_void_p] * 8),
    'salsa_fasta_headers': (C.c_int, [c_void_p, c_size_t, c_void_p, c_void_p, C.c_uint32, c_void_p]),
    'salsa_fasta_tiles': (c_size_t, [c_size_t]),
    'salsa_fasta_count': (C.c_int, [c_void_p, c_size_t, c_void_p, c_void_p, c_i32, c_void_p, c_void_p, c_void_p]),
    'salsa_exclusive_scan_u64': (C.c_int, [c_void_p, c_size_t, c_void_p]),
    'salsa_break_contigs_run': (C.c_int, [c_i32, c_i32, c_void_p, c_void_p, c_size_t, C.c_char_p]),
    'salsa_reproject_candidates': (C.c_int, [c_void_p, c_void_p, c_size_t, c_i32, c_void_p, c_void_p, c_size_t, c_void_p,
                                             c_void_p, c_void_p]),
    'salsa_sort_u64': (C.c_int, [c_void_p, c_void_p, c_size_t, C.c_uint64, c_void_p, C.POINTER(c_size_t), c_void_p]),
    'salsa_reproject_blocks': (c_size_t, [c_size_t]),
    'salsa_reproject_emit': (C.c_int, [c_void_p, c_void_p, c_size_t, c_i32, c_void_p, c_size_t] + [c_void_p] * 9 +
                             [c_size_t, c_void_p, c_void_p]),
    'salsa_fasta_span_lengths': (C.c_int, [c_void_p, c_size_t, c_void_p, c_void_p, c_i32, c_void_p, c_void_p, c_void_p,
                                           c_void_p, c_void_p]),
    'salsa_fasta_pack': (C.c_int, [c_void_p, c_size_t, c_void_p, c_void_p, c_i32, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_void_p]),
    'salsa_re_count': (C.c_int, [c_void_p, c_void_p, c_void_p, c_i32, c_u64, C.c_char_p, c_i32, c_void_p, c_void_p,
                                 c_void_p, c_void_p]),
    'salsa_cov_scan_start': (C.c_int, [c_void_p, c_void_p, c_size_t, c_u64, c_i32, c_void_p, c_void_p, c_void_p, c_void_p,
                                       c_void_p, c_void_p]),
    'salsa_cov_scan_joins': (C.c_int, [c_void_p, c_void_p, c_size_t, c_u64, c_i32, c_void_p, c_void_p, c_void_p, c_void_p,
                                       c_void_p, c_void_p, c_void_p, c_void_p]),
    'salsa_cov_scratch_bytes': (c_size_t, [c_void_p, c_i32]),
    'salsa_cov_analyse': (C.c_int, [c_void_p, c_void_p, c_void_p, c_i32, c_void_p, c_size_t, c_void_p, c_void_p, c_void_p,
                                    c_void_p]),
    'salsa_format_g12': (C.c_int, [c_void_p, c_void_p, c_size_t, c_void_p, c_size_t, c_void_p]),
    'salsa_format_links': (c_i64, [C.c_char_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_size_t, c_void_p, c_size_t]),
    'salsa_format_scaled': (c_i64, [C.c_char_p] + [c_void_p] * 11 + [c_size_t, c_void_p, c_size_t]),
}


def lib():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError('libsalsa_b200.so not built (%s); run `make -C salsa_b200/csrc` or '
                               '__graft_entry__.build() -- there is no CPU fallback' % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise SalsaError(rc, lib().salsa_last_error().decode('latin-1'))
