"""ctypes mirror of include/hj.h (the C ABI of libhj). Field order and widths must match the header exactly;
tests/test_abi.py checks the struct sizes against the values compiled into the library (hj_abi_sizeof)."""
import ctypes as C

HJ_ABI_VERSION = 3

# JoinMode, src/lib/types.hpp:189 of the reference
MODE_INNER, MODE_LEFT, MODE_RIGHT, MODE_OUTER, MODE_CROSS, MODE_SEMI, MODE_ANTI = range(7)
MODES = {"Inner": MODE_INNER, "Left": MODE_LEFT, "Right": MODE_RIGHT, "Outer": MODE_OUTER, "Cross": MODE_CROSS,
         "Semi": MODE_SEMI, "Anti": MODE_ANTI}
KEY_INT32, KEY_INT64, KEY_FLOAT, KEY_DOUBLE = 0, 1, 2, 3
SEG_VALUE, SEG_DICT, SEG_GATHERED, SEG_REFERENCE, SEG_FOR, SEG_RLE = 0, 1, 2, 3, 4, 5
LOC_HOST, LOC_DEVICE = 0, 1
FLAG_SQL_NULL_PROBE = 0x1
FLAG_TIMINGS = 0x2
FLAG_XCHG_STAGED = 0x4
FLAG_TWO_SWEEP_PARTITION = 0x8
FLAG_XCHG_GENERAL = 0x10
INVALID_CHUNK_ID = 0xFFFFFFFF
INVALID_CHUNK_OFFSET = 0xFFFFFFFF
N_PHASES = 10
PHASES = ["h2d", "materialize", "scatter1", "scatter2", "build_probe", "finalize", "d2h", "total", "xchg_partition",
          "xchg_transfer"]
NCCL_UNIQUE_ID_BYTES = 128

OK, ERR_INVALID, ERR_CUDA, ERR_OOM, ERR_CAPACITY, ERR_TOO_LARGE, ERR_STATE, ERR_COMM = range(8)


class RowID(C.Structure):
    _fields_ = [("chunk_id", C.c_uint32), ("chunk_offset", C.c_uint32)]


class Segment(C.Structure):
    _fields_ = [
        ("chunk_id", C.c_uint32), ("n_rows", C.c_uint32),
        ("kind", C.c_uint8), ("key_type", C.c_uint8), ("attr_width", C.c_uint8), ("reserved0", C.c_uint8),
        ("dict_n", C.c_uint32), ("null_value_id", C.c_uint32), ("reserved1", C.c_uint32),
        ("values", C.c_void_p), ("nulls", C.c_void_p), ("attr", C.c_void_p), ("row_offsets", C.c_void_p),
    ]


class Relation(C.Structure):
    _fields_ = [("segs", C.POINTER(Segment)), ("n_segs", C.c_uint32), ("location", C.c_uint32),
                ("n_rows", C.c_uint64)]


class Options(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("flags", C.c_uint32), ("radix_bits", C.c_int32),
                ("reserved0", C.c_uint32), ("output_capacity", C.c_uint64), ("stream_slice_rows", C.c_uint64),
                ("reserved", C.c_uint64 * 3)]


class ResultInfo(C.Structure):
    _fields_ = [("n_rows", C.c_uint64), ("n_sides", C.c_uint32), ("radix_bits", C.c_uint32),
                ("n_build_tuples", C.c_uint64), ("n_probe_tuples", C.c_uint64),
                ("n_partitions", C.c_uint32), ("n_overflow_partitions", C.c_uint32),
                ("n_probe_passes", C.c_uint32), ("n_probe_slices", C.c_uint32)]


class Aggregate(C.Structure):
    _fields_ = [("n_rows", C.c_uint64), ("n_values", C.c_uint64), ("is_floating", C.c_int32), ("reserved", C.c_int32),
                ("sum_int", C.c_int64), ("sum_double", C.c_double), ("min_int", C.c_int64), ("max_int", C.c_int64),
                ("min_double", C.c_double), ("max_double", C.c_double)]


class Group(C.Structure):
    _fields_ = [("key", C.c_int64), ("key_is_null", C.c_int32), ("reserved", C.c_int32), ("agg", Aggregate)]


class GroupColumn(C.Structure):
    _fields_ = [("values", C.c_void_p), ("nulls", C.c_void_p), ("key_type", C.c_int32), ("reserved", C.c_int32)]


class GroupInfo(C.Structure):
    _fields_ = [("n_groups", C.c_uint64), ("n_rows", C.c_uint64)]


class ScanInfo(C.Structure):
    _fields_ = [("n_rows", C.c_uint64), ("n_chunks", C.c_uint32), ("n_nonempty_chunks", C.c_uint32)]


# PredicateCondition, src/lib/types.hpp:166-179 of the reference
PRED_EQUALS, PRED_NOT_EQUALS, PRED_LESS_THAN, PRED_LESS_THAN_EQUALS, PRED_GREATER_THAN, PRED_GREATER_THAN_EQUALS, \
    PRED_BETWEEN, PRED_IN, PRED_LIKE, PRED_NOT_LIKE, PRED_IS_NULL, PRED_IS_NOT_NULL = range(12)
PREDICATES = {"Equals": PRED_EQUALS, "NotEquals": PRED_NOT_EQUALS, "LessThan": PRED_LESS_THAN,
              "LessThanEquals": PRED_LESS_THAN_EQUALS, "GreaterThan": PRED_GREATER_THAN,
              "GreaterThanEquals": PRED_GREATER_THAN_EQUALS, "Between": PRED_BETWEEN, "In": PRED_IN, "Like": PRED_LIKE,
              "NotLike": PRED_NOT_LIKE, "IsNull": PRED_IS_NULL, "IsNotNull": PRED_IS_NOT_NULL}
SCAN_SLOTS = 4
SCAN_DEREFERENCE = 0x1
SCAN_VALUE_IS_DOUBLE = 0x2


class Timings(C.Structure):
    _fields_ = [("ms", C.c_float * N_PHASES), ("bytes_h2d", C.c_uint64), ("bytes_d2h", C.c_uint64),
                ("n_kernel_launches", C.c_uint32), ("xchg_path", C.c_uint32), ("bytes_nvlink_sent", C.c_uint64)]


# every symbol include/hj.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "hj_abi_version": (C.c_int, []),
    "hj_abi_sizeof": (C.c_size_t, [C.c_int]),
    "hj_create": (C.c_int, [C.c_int, C.POINTER(Options), C.POINTER(C.c_void_p)]),
    "hj_destroy": (None, [C.c_void_p]),
    "hj_last_error": (C.c_char_p, [C.c_void_p]),
    "hj_radix_bits": (C.c_int, [C.c_uint64, C.c_int]),
    "hj_join": (C.c_int, [C.c_void_p, C.POINTER(Relation), C.POINTER(Relation), C.c_int, C.c_void_p,
                          C.POINTER(ResultInfo)]),
    "hj_fetch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]),
    "hj_bind_output": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]),
    "hj_result_device": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]),
    "hj_translate": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint32),
                               C.c_uint32]),
    "hj_get_timings": (C.c_int, [C.c_void_p, C.POINTER(Timings)]),
    "hj_release_scratch": (C.c_int, [C.c_void_p]),
    "hj_scan": (C.c_int, [C.c_void_p, C.POINTER(Relation), C.c_int, C.c_int64, C.c_int, C.c_int, C.c_uint32, C.c_void_p,
                          C.POINTER(ScanInfo)]),
    "hj_scan_value_ids": (C.c_int, [C.c_void_p, C.POINTER(Relation), C.c_int, C.c_void_p, C.c_int, C.c_uint32, C.c_void_p,
                                    C.POINTER(ScanInfo)]),
    "hj_translate_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(Relation), C.c_void_p]),
    "hj_gather_column": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(Relation), C.c_int, C.c_void_p, C.POINTER(C.c_void_p),
                                   C.POINTER(C.c_void_p), C.POINTER(C.c_int)]),
    "hj_aggregate_column": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.POINTER(Aggregate)]),
    "hj_aggregate_column_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]),
    "hj_aggregate_grouped": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_uint64,
                                       C.c_uint64, C.c_void_p, C.POINTER(GroupInfo)]),
    "hj_aggregate_grouped_rows": (C.c_int, [C.c_void_p, C.POINTER(GroupColumn), C.c_uint32, C.c_void_p, C.c_void_p, C.c_int, C.c_uint64,
                                            C.c_uint64, C.c_void_p, C.POINTER(GroupInfo)]),
    "hj_aggregate_grouped_rows_relation": (C.c_int, [C.c_void_p, C.POINTER(C.POINTER(Relation)), C.c_uint32, C.POINTER(Relation),
                                                     C.c_uint64, C.c_void_p, C.POINTER(GroupInfo)]),
    "hj_aggregate_grouped_relation": (C.c_int, [C.c_void_p, C.POINTER(Relation), C.POINTER(Relation), C.c_uint64, C.c_void_p,
                                                C.POINTER(GroupInfo)]),
    "hj_aggregate_grouped_fetch": (C.c_int, [C.c_void_p, C.POINTER(Group), C.c_uint64]),
    "hj_aggregate_relation": (C.c_int, [C.c_void_p, C.POINTER(Relation), C.c_void_p, C.POINTER(Aggregate)]),
    "hj_scan_result_device": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                        C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]),
    "hj_scan_fetch": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p]),
    "hj_scan_release": (C.c_int, [C.c_void_p, C.c_int]),
    "hj_exchange_rank_of": (C.c_uint32, [C.c_int64, C.c_int, C.c_uint32]),
    "hj_comm_unique_id": (C.c_int, [C.POINTER(C.c_uint8)]),
    "hj_comm_init": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint8), C.c_int, C.c_int]),
    "hj_join_sharded": (C.c_int, [C.c_void_p, C.POINTER(Relation), C.POINTER(Relation), C.c_int, C.c_void_p,
                                  C.POINTER(ResultInfo)]),
}


def bind(lib):
    """Sets restype/argtypes for every declared symbol; raises AttributeError if one is missing."""
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib
