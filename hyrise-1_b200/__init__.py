"""hyrise-1_b200 — B200-native JoinHash (Hyrise's radix-partitioned hash equi-join) behind a C ABI.

The product is `csrc/libhj.so` (hand-written sm_100a CUDA kernels + the C ABI of include/hj.h) and the C++ host
shim in `host/` that keeps the reference's `JoinHash : AbstractJoinOperator` interface. This Python module is only
the thin ctypes binding the tests and bench.py drive the library through; it contains no join logic and no CPU
fallback: loading fails loudly when the CUDA library has not been built.

Import with `importlib.import_module("hyrise-1_b200")` (the directory name is not a Python identifier).
"""
import ctypes as C
import importlib
import os

import numpy as np

abi = importlib.import_module(__name__ + ".abi")

_HERE = os.path.dirname(os.path.abspath(__file__))
# HJ_LIB_VARIANT=name loads csrc/variants/libhj_<name>.so instead (tuning builds of tests/build_variants.sh; development only)
LIB_PATH = (os.path.join(_HERE, "csrc", "variants", f"libhj_{os.environ['HJ_LIB_VARIANT']}.so") if os.environ.get("HJ_LIB_VARIANT")
            else os.path.join(_HERE, "csrc", "libhj.so"))
_lib = None


class HjError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"libhj status {status}: {msg}")
        self.status = status


def lib():
    """The loaded libhj.so with typed entry points. Raises if the extension is missing — there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'`")
        _lib = abi.bind(C.CDLL(LIB_PATH))
        if _lib.hj_abi_version() != abi.HJ_ABI_VERSION:
            raise ImportError("libhj.so ABI version does not match hyrise-1_b200/abi.py")
    return _lib


def _ptr_and_keep(x):
    """(address, object to keep alive) for a numpy array or a torch tensor (host or device)."""
    if x is None:
        return None, None
    if isinstance(x, np.ndarray):
        assert x.flags["C_CONTIGUOUS"]
        return x.ctypes.data, x
    # torch tensor (duck-typed so that torch is not imported when only numpy is used)
    assert x.is_contiguous()
    return x.data_ptr(), x


_KEY_DTYPES = {"int32": abi.KEY_INT32, "int64": abi.KEY_INT64, "float32": abi.KEY_FLOAT, "float64": abi.KEY_DOUBLE}


def _dtype_name(x):
    return str(x.dtype).replace("torch.", "")


class RelationBuilder:
    """Collects the per-chunk segment descriptors of one key column (hj_relation)."""

    def __init__(self, location=abi.LOC_HOST):
        self.location = location
        self._segs = []
        self._keep = []

    def _add(self, seg, *keep):
        self._segs.append(seg)
        self._keep.extend(keep)
        return self

    def add_value(self, chunk_id, values, nulls=None, kind=abi.SEG_VALUE, row_offsets=None):
        """ValueColumn<T> (or a pre-gathered ReferenceColumn with kind=SEG_GATHERED)."""
        s = abi.Segment()
        s.chunk_id, s.n_rows, s.kind = chunk_id, len(values), kind
        s.key_type = _KEY_DTYPES[_dtype_name(values)]
        s.values, k1 = _ptr_and_keep(values)
        s.nulls, k2 = _ptr_and_keep(nulls)
        s.row_offsets, k3 = _ptr_and_keep(row_offsets)
        if nulls is not None:
            assert _dtype_name(nulls) in ("uint8", "bool") and len(nulls) == len(values)
        if row_offsets is not None:
            assert _dtype_name(row_offsets) in ("uint32", "int32") and len(row_offsets) == len(values)
        return self._add(s, k1, k2, k3)

    def add_dict(self, chunk_id, dictionary, attr, null_value_id=None):
        """DictionaryColumn<T> with a FixedSizeByteAligned attribute vector (uint8/uint16/uint32 or torch int types)."""
        s = abi.Segment()
        s.chunk_id, s.n_rows, s.kind = chunk_id, len(attr), abi.SEG_DICT
        s.key_type = _KEY_DTYPES[_dtype_name(dictionary)]
        width = {"uint8": 1, "uint16": 2, "int16": 2, "uint32": 4, "int32": 4}[_dtype_name(attr)]
        s.attr_width = width
        s.dict_n = len(dictionary)
        s.null_value_id = len(dictionary) if null_value_id is None else null_value_id
        s.values, k1 = _ptr_and_keep(dictionary)
        s.attr, k2 = _ptr_and_keep(attr)
        return self._add(s, k1, k2)

    def add_value_ids(self, chunk_id, attr, null_value_id):
        """A dictionary chunk of any value type for hj_scan_value_ids: the attribute vector and the NULL value id only —
        the dictionary stays with the caller."""
        s = abi.Segment()
        s.chunk_id, s.n_rows, s.kind = chunk_id, len(attr), abi.SEG_DICT
        s.key_type = abi.KEY_INT32
        s.attr_width = {"uint8": 1, "uint16": 2, "int16": 2, "uint32": 4, "int32": 4}[_dtype_name(attr)]
        s.dict_n = 0
        s.null_value_id = null_value_id
        s.attr, k = _ptr_and_keep(attr)
        return self._add(s, k)

    def add_dict_bp128(self, chunk_id, dictionary, stream, n_rows, null_value_id=None):
        """DictionaryColumn<T> whose attribute vector is SimdBp128-compressed: `stream` = the uint128 words as uint32[4*words]."""
        s = abi.Segment()
        s.chunk_id, s.n_rows, s.kind = chunk_id, n_rows, abi.SEG_DICT
        s.key_type = _KEY_DTYPES[_dtype_name(dictionary)]
        s.attr_width = 0
        s.dict_n = len(dictionary)
        s.null_value_id = len(dictionary) if null_value_id is None else null_value_id
        s.values, k1 = _ptr_and_keep(dictionary)
        s.attr, k2 = _ptr_and_keep(stream)
        return self._add(s, k1, k2)

    def add_for_bp128(self, chunk_id, block_minima, stream, n_rows, nulls=None):
        """FrameOfReferenceColumn<T> with SimdBp128-compressed offsets."""
        s = abi.Segment()
        s.chunk_id, s.n_rows, s.kind = chunk_id, n_rows, abi.SEG_FOR
        s.key_type = _KEY_DTYPES[_dtype_name(block_minima)]
        s.attr_width = 0
        s.values, k1 = _ptr_and_keep(block_minima)
        s.attr, k2 = _ptr_and_keep(stream)
        s.nulls, k3 = _ptr_and_keep(nulls)
        return self._add(s, k1, k2, k3)

    def add_for(self, chunk_id, block_minima, offsets, nulls=None):
        """FrameOfReferenceColumn<T>: minima T[ceil(n/2048)], FixedSizeByteAligned offsets uint8/16/32[n], 1 byte/row flags."""
        s = abi.Segment()
        s.chunk_id, s.n_rows, s.kind = chunk_id, len(offsets), abi.SEG_FOR
        s.key_type = _KEY_DTYPES[_dtype_name(block_minima)]
        s.attr_width = {"uint8": 1, "uint16": 2, "int16": 2, "uint32": 4, "int32": 4}[_dtype_name(offsets)]
        assert len(block_minima) == (len(offsets) + 2047) // 2048
        s.values, k1 = _ptr_and_keep(block_minima)
        s.attr, k2 = _ptr_and_keep(offsets)
        s.nulls, k3 = _ptr_and_keep(nulls)
        return self._add(s, k1, k2, k3)

    def add_rle(self, chunk_id, n_rows, run_values, end_positions, run_nulls=None):
        """RunLengthColumn<T>: run values T[runs], inclusive end positions uint32[runs], 1 byte per run flags."""
        s = abi.Segment()
        s.chunk_id, s.n_rows, s.kind = chunk_id, n_rows, abi.SEG_RLE
        s.key_type = _KEY_DTYPES[_dtype_name(run_values)]
        s.dict_n = len(run_values)
        assert len(end_positions) == len(run_values) and _dtype_name(end_positions) in ("uint32", "int32")
        s.values, k1 = _ptr_and_keep(run_values)
        s.attr, k2 = _ptr_and_keep(end_positions)
        s.nulls, k3 = _ptr_and_keep(run_nulls)
        return self._add(s, k1, k2, k3)

    def add_reference(self, chunk_id, pos_list, referenced, key_type):
        """ReferenceColumn as stored: pos_list = (n, 2) uint32 [chunk_id, chunk_offset] rows (NULL_ROW_ID = NULL key) into
        `referenced`, a built relation describing the key column of the data table (same location). The library gathers
        the keys on the device (HJ_SEG_REFERENCE)."""
        s = abi.Segment()
        s.chunk_id, s.n_rows, s.kind = chunk_id, len(pos_list), abi.SEG_REFERENCE
        s.key_type = key_type
        s.attr, k1 = _ptr_and_keep(pos_list)
        s.values = C.cast(C.pointer(referenced), C.c_void_p)
        return self._add(s, k1, referenced)

    def build(self):
        n = len(self._segs)
        arr = (abi.Segment * max(n, 1))(*self._segs)
        rel = abi.Relation()
        rel.segs = C.cast(arr, C.POINTER(abi.Segment))
        rel.n_segs = n
        rel.location = self.location
        rel.n_rows = sum(s.n_rows for s in self._segs)
        rel._keep = (arr, self._keep)  # keep the descriptor array and the data alive
        return rel


def rowids_to_numpy(buf, n):
    """View n hj_row_id structs as an (n, 2) uint32 array [chunk_id, chunk_offset]."""
    return np.frombuffer(buf, dtype=np.uint32, count=2 * n).reshape(n, 2)


class HashJoin:
    """One hj_ctx. Mirrors the call sequence of the C++ shim's JoinHash::_on_execute: join -> fetch."""

    def __init__(self, device=0, flags=0, radix_bits=-1, output_capacity=0, stream_slice_rows=0):
        self._l = lib()
        opts = abi.Options()
        opts.abi_version = abi.HJ_ABI_VERSION
        opts.flags = flags
        opts.radix_bits = radix_bits
        opts.output_capacity = output_capacity
        opts.stream_slice_rows = stream_slice_rows
        self._ctx = C.c_void_p()
        st = self._l.hj_create(device, C.byref(opts), C.byref(self._ctx))
        if st != abi.OK:
            raise HjError(st, (self._l.hj_last_error(None) or b"").decode())
        self.info = None
        self._mode = None

    def close(self):
        if getattr(self, "_ctx", None):
            self._l.hj_destroy(self._ctx)
            self._ctx = None

    __del__ = close

    def _check(self, st):
        if st != abi.OK:
            raise HjError(st, (self._l.hj_last_error(self._ctx) or b"").decode())

    def join(self, build, probe, mode, stream=None):
        """build/probe: abi.Relation after side selection; mode: abi.MODE_* or its name. Returns hj_result_info."""
        if isinstance(mode, str):
            mode = abi.MODES[mode]
        info = abi.ResultInfo()
        self.info = None
        self._check(self._l.hj_join(self._ctx, C.byref(build), C.byref(probe), mode, stream, C.byref(info)))
        self.info, self._mode = info, mode
        return info

    def join_sharded(self, build, probe, mode, stream=None):
        if isinstance(mode, str):
            mode = abi.MODES[mode]
        info = abi.ResultInfo()
        self._check(self._l.hj_join_sharded(self._ctx, C.byref(build), C.byref(probe), mode, stream, C.byref(info)))
        self.info, self._mode = info, mode
        return info

    def bind_output(self, out_build, out_probe):
        """hj_bind_output: the NEXT join writes its PosLists straight into these host arrays ((cap, 2) uint32, ideally
        pinned); with host inputs of many chunks the join is streamed (copy-in, join and copy-out overlap).
        Afterwards fetch(out_build, out_probe) just returns the filled views."""
        cap = len(out_probe)
        if out_build is not None:
            cap = min(cap, len(out_build))
        self._bound = (out_build, out_probe)  # keep the arrays alive until the join has run
        self._check(self._l.hj_bind_output(self._ctx, None if out_build is None else out_build.ctypes.data,
                                           out_probe.ctypes.data, cap))

    def fetch(self, out_build=None, out_probe=None):
        """Copies the result PosLists to host. Returns (build_side or None, probe_side) as (n, 2) uint32 arrays."""
        if self.info is None:
            self._check(self._l.hj_fetch(self._ctx, None, None, 0))  # reports HJ_ERR_STATE
        n = self.info.n_rows
        two = self.info.n_sides == 2
        if out_probe is None:
            out_probe = np.empty((max(n, 1), 2), dtype=np.uint32)
        if two and out_build is None:
            out_build = np.empty((max(n, 1), 2), dtype=np.uint32)
        self._check(self._l.hj_fetch(self._ctx, out_build.ctypes.data if two else None, out_probe.ctypes.data, n))
        return (out_build[:n] if two else None), out_probe[:n]

    def result_device(self):
        """(build_side device ptr or None, probe_side device ptr, n_rows) of the last join."""
        b, p, n = C.c_void_p(), C.c_void_p(), C.c_uint64()
        self._check(self._l.hj_result_device(self._ctx, C.byref(b), C.byref(p), C.byref(n)))
        return b.value, p.value, n.value

    def translate(self, rowids, pos_lists):
        """hj_translate: rowids (n,2) uint32 in place through per-chunk pos lists ((m,2) uint32 arrays)."""
        n_chunks = len(pos_lists)
        ptrs = (C.c_void_p * max(n_chunks, 1))(*[p.ctypes.data for p in pos_lists])
        sizes = (C.c_uint32 * max(n_chunks, 1))(*[len(p) for p in pos_lists])
        self._check(self._l.hj_translate(self._ctx, rowids.ctypes.data, len(rowids), ptrs, sizes, n_chunks))
        return rowids

    # ---- TableScan on the device -------------------------------------------------------------------
    def scan(self, column, predicate, value=None, slot=0, dereference=False, stream=None):
        """hj_scan: `column <predicate> value` (value None = NULL literal; a Python / numpy float is passed as a double
        literal, HJ_SCAN_VALUE_IS_DOUBLE). Returns hj_scan_info; the PosLists stay on the device in `slot`."""
        if isinstance(predicate, str):
            predicate = abi.PREDICATES[predicate]
        info = abi.ScanInfo()
        flags = abi.SCAN_DEREFERENCE if dereference else 0
        if value is None:
            bits = 0
        elif isinstance(value, (float, np.floating)):
            bits = int(np.array([value], dtype=np.float64).view(np.int64)[0])
            flags |= abi.SCAN_VALUE_IS_DOUBLE
        else:
            bits = int(value)
        self._check(self._l.hj_scan(self._ctx, C.byref(column), predicate, bits, 1 if value is None else 0, slot, flags, stream,
                                    C.byref(info)))
        return info

    def scan_value_ids(self, column, predicate, bounds=None, slot=0, stream=None):
        """hj_scan_value_ids: dictionary chunks of any value type scanned on their value ids; bounds = (n_chunks, 2) uint32
        [lower bound, upper bound] of the literal in every chunk's dictionary, found by the caller."""
        if isinstance(predicate, str):
            predicate = abi.PREDICATES[predicate]
        info = abi.ScanInfo()
        if bounds is not None:
            bounds = np.ascontiguousarray(bounds, dtype=np.uint32)
            assert bounds.size == 2 * column.n_segs
        self._check(self._l.hj_scan_value_ids(self._ctx, C.byref(column), predicate, None if bounds is None else bounds.ctypes.data,
                                              slot, 0, stream, C.byref(info)))
        return info

    def translate_device(self, device_inout, n_rows, input_relation, stream=None):
        """hj_translate_device: RowIDs of a device-resident reference-table input -> rows of the referenced data table, in place."""
        self._check(self._l.hj_translate_device(self._ctx, device_inout, n_rows, C.byref(input_relation), stream))

    def gather_column(self, device_pos, n_rows, column, slot=0, stream=None):
        """hj_gather_column: (values device ptr, nulls device ptr, key type) of `column` at the rows of a device PosList;
        the output stays valid until the next gather into the same slot."""
        v, n, t = C.c_void_p(), C.c_void_p(), C.c_int()
        self._check(self._l.hj_gather_column(self._ctx, device_pos, n_rows, C.byref(column), slot, stream, C.byref(v), C.byref(n),
                                             C.byref(t)))
        return v.value, n.value, t.value

    def aggregate_column(self, device_values, device_nulls, n_rows, key_type, stream=None):
        """hj_aggregate_column: COUNT / SUM / MIN / MAX (and AVG = sum / count) of a device column, as a dict with the
        reference's result types; SUM / MIN / MAX / AVG are None when every row is NULL (aggregate.cpp:635-642)."""
        a = abi.Aggregate()
        self._check(self._l.hj_aggregate_column(self._ctx, device_values, device_nulls, n_rows, key_type, stream, C.byref(a)))
        return self._aggregate_dict(a)

    def aggregate_column_static(self, device_values, device_nulls, n_rows, key_type, device_out, stream=None):
        """hj_aggregate_column_static: the same aggregates without any allocation, synchronisation or host copy; the result
        is one hj_aggregate record (abi.Aggregate, 80 bytes) written to the DEVICE address `device_out` in stream order.
        `aggregate_record(bytes)` turns a copy of that record into the dict aggregate_column returns."""
        self._check(self._l.hj_aggregate_column_static(self._ctx, device_values, device_nulls, n_rows, key_type, stream, device_out))

    @classmethod
    def aggregate_record(cls, raw):
        return cls._aggregate_dict(abi.Aggregate.from_buffer_copy(bytes(raw)))

    def aggregate_grouped(self, group_keys, group_nulls, group_key_type, values, value_nulls, value_type, n_rows, max_groups=0,
                          stream=None):
        """hj_aggregate_grouped + fetch: {group key or None: aggregate dict} for device columns (pointers as ints)."""
        info = abi.GroupInfo()
        self._check(self._l.hj_aggregate_grouped(self._ctx, group_keys, group_nulls, group_key_type, values, value_nulls, value_type,
                                                 n_rows, max_groups, stream, C.byref(info)))
        return self._fetch_groups(info)

    def aggregate_grouped_rows(self, group_columns, values, value_nulls, value_type, n_rows, max_groups=0, stream=None):
        """hj_aggregate_grouped_rows + fetch: group_columns = [(values ptr, nulls ptr or None, key type)]; returns
        {representative row index: aggregate dict}."""
        cols = (abi.GroupColumn * len(group_columns))()
        for c, (v, z, kt) in zip(cols, group_columns):
            c.values, c.nulls, c.key_type = v, z, kt
        info = abi.GroupInfo()
        self._check(self._l.hj_aggregate_grouped_rows(self._ctx, cols, len(group_columns), values, value_nulls, value_type, n_rows,
                                                      max_groups, stream, C.byref(info)))
        return self._fetch_groups(info)

    def aggregate_grouped_rows_relation(self, group_columns, value_column=None, max_groups=0, stream=None):
        """hj_aggregate_grouped_rows_relation + fetch: {representative row number in the table: aggregate dict}."""
        ptrs = (C.POINTER(abi.Relation) * len(group_columns))(*[C.pointer(r) for r in group_columns])
        info = abi.GroupInfo()
        self._check(self._l.hj_aggregate_grouped_rows_relation(self._ctx, ptrs, len(group_columns),
                                                               None if value_column is None else C.byref(value_column), max_groups,
                                                               stream, C.byref(info)))
        return self._fetch_groups(info)

    def aggregate_grouped_relation(self, group_column, value_column=None, max_groups=0, stream=None):
        """hj_aggregate_grouped_relation + fetch: both columns as chunk descriptors of one table (value_column None: COUNT(*))."""
        info = abi.GroupInfo()
        self._check(self._l.hj_aggregate_grouped_relation(self._ctx, C.byref(group_column),
                                                          None if value_column is None else C.byref(value_column), max_groups, stream,
                                                          C.byref(info)))
        return self._fetch_groups(info)

    def _fetch_groups(self, info):
        groups = (abi.Group * max(info.n_groups, 1))()
        self._check(self._l.hj_aggregate_grouped_fetch(self._ctx, groups, info.n_groups))
        out = {}
        for i in range(info.n_groups):
            g = groups[i]
            key = None if g.key_is_null else g.key
            assert key not in out, "a group was emitted twice"
            out[key] = self._aggregate_dict(g.agg)
        return out

    def aggregate_relation(self, column, stream=None):
        """hj_aggregate_relation: the same aggregates over a column given as chunk descriptors (host or device resident,
        data or reference table)."""
        a = abi.Aggregate()
        self._check(self._l.hj_aggregate_relation(self._ctx, C.byref(column), stream, C.byref(a)))
        return self._aggregate_dict(a)

    @staticmethod
    def _aggregate_dict(a):
        some = a.n_values > 0
        floating = bool(a.is_floating)
        return {"count_rows": a.n_rows, "count": a.n_values,
                "sum": None if not some else (a.sum_double if floating else a.sum_int),
                "min": None if not some else (a.min_double if floating else a.min_int),
                "max": None if not some else (a.max_double if floating else a.max_int),
                "avg": None if not some else a.sum_double / a.n_values}

    def scan_fetch(self, slot=0):
        """(pos (n, 2) uint32 [chunk_id, chunk_offset], chunk_begin (n_chunks + 1,) uint32) of the scan result in `slot`."""
        _, _, n, n_chunks = self.scan_result_device(slot)
        pos = np.empty((max(n, 1), 2), dtype=np.uint32)
        begin = np.empty(n_chunks + 1, dtype=np.uint32)
        self._check(self._l.hj_scan_fetch(self._ctx, slot, pos.ctypes.data, n, begin.ctypes.data))
        return pos[:n], begin

    def scan_chunk_begin(self, slot=0):
        """chunk_begin (n_chunks + 1,) uint32 of the scan result in `slot` without downloading the PosLists."""
        _, _, _, n_chunks = self.scan_result_device(slot)
        begin = np.empty(n_chunks + 1, dtype=np.uint32)
        self._check(self._l.hj_scan_fetch(self._ctx, slot, None, 0, begin.ctypes.data))
        return begin

    def scan_result_device(self, slot=0):
        """(pos device ptr, chunk_begin device ptr, n_rows, n_chunks) of the scan result in `slot`."""
        p, b, n, c = C.c_void_p(), C.c_void_p(), C.c_uint64(), C.c_uint32()
        self._check(self._l.hj_scan_result_device(self._ctx, slot, C.byref(p), C.byref(b), C.byref(n), C.byref(c)))
        return p.value, b.value, n.value, c.value

    def scan_release(self, slot=-1):
        self._check(self._l.hj_scan_release(self._ctx, slot))

    def timings(self):
        t = abi.Timings()
        self._check(self._l.hj_get_timings(self._ctx, C.byref(t)))
        d = {name: t.ms[i] for i, name in enumerate(abi.PHASES)}
        d.update(bytes_h2d=t.bytes_h2d, bytes_d2h=t.bytes_d2h, n_kernel_launches=t.n_kernel_launches,
                 bytes_nvlink_sent=t.bytes_nvlink_sent, xchg_path=t.xchg_path)
        return d

    def release_scratch(self):
        self._check(self._l.hj_release_scratch(self._ctx))

    def comm_init(self, unique_id, n_ranks, rank):
        buf = (C.c_uint8 * abi.NCCL_UNIQUE_ID_BYTES)(*unique_id)
        self._check(self._l.hj_comm_init(self._ctx, buf, n_ranks, rank))


def comm_unique_id():
    buf = (C.c_uint8 * abi.NCCL_UNIQUE_ID_BYTES)()
    st = lib().hj_comm_unique_id(buf)
    if st != abi.OK:
        raise HjError(st, "hj_comm_unique_id failed")
    return bytes(buf)


def exchange_rank_of(key, key_type, n_ranks):
    """Rank on which all rows with `key` meet in hj_join_sharded (host arithmetic, no GPU needed)."""
    return lib().hj_exchange_rank_of(int(key), key_type, n_ranks)


def radix_bits(n_build_rows, key_type=abi.KEY_INT32):
    return lib().hj_radix_bits(n_build_rows, key_type)
