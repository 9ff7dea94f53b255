"""CPU restatement of the reference's single-column TableScan for int / long / float / double columns, and for string
columns compared with <, <=, =, !=, >=, > (numpy). TEST INFRASTRUCTURE: imported
only by tests/ — never by the product package (hyrise-1_b200 has no CPU path).

Follows, per chunk, what TableScan::_on_execute dispatches to (src/lib/operators/table_scan.cpp:84-171, _init_scan :173-201):
  * SingleColumnTableScanImpl (operators/table_scan/single_column_table_scan_impl.cpp)
      - NULL literal -> empty PosList (:25-34)
      - ValueColumn / other encoded columns: NULL rows skipped, comparator on the value
        (:41-81, BaseTableScanImpl::_unary_scan_with_value, base_table_scan_impl.hpp:52-65)
      - DictionaryColumn: the literal's lower/upper bound in the sorted dictionary (_get_search_value_id :135-150), the
        "matches all"/"matches none" early outs (:152-199) and otherwise a comparison of VALUE IDS (:83-133,
        table of conditions :93-99)
  * IsNullTableScanImpl (is_null_table_scan_impl.cpp:24-133): IS [NOT] NULL on null flags / null value ids, and for a
    ReferenceColumn additionally the NULL RowIDs of its own PosList (:24-37)
  * BaseSingleColumnTableScanImpl::handle_column(ReferenceColumn) (base_single_column_table_scan_impl.cpp:33-60): a
    reference chunk is scanned through its PosList; matches are offsets INTO THE REFERENCE CHUNK.
A chunk's result is the ascending list of matching chunk offsets; the operator turns them into RowID{chunk_id, offset}
(data table) or pos_list[offset] (reference table, table_scan.cpp:112-131).

Chunk descriptions (plain dicts of numpy arrays):
  {"kind": "value", "values": int array, "nulls": uint8 array or None}
  {"kind": "dict", "dictionary": sorted unique int array, "attr": uint array, "null_value_id": int}
  {"kind": "for", "minima": int array (one per 2048 rows), "offsets": uint array, "nulls": uint8 array or None}
  {"kind": "rle", "run_values": int array, "end_positions": uint32 array (inclusive), "run_nulls": uint8 array or None}
      (both scanned as decoded values: handle_column(BaseEncodedColumn), single_column_table_scan_impl.cpp:58-81)
  {"kind": "reference", "pos_list": (n, 2) uint32 array [chunk_id, chunk_offset] (0xFFFFFFFF pairs = NULL_ROW_ID),
   "referenced": [chunk descriptions of the referenced data table's column]}
"""
import numpy as np

EQUALS, NOT_EQUALS, LESS_THAN, LESS_THAN_EQUALS, GREATER_THAN, GREATER_THAN_EQUALS = range(6)  # types.hpp:166-172
IS_NULL, IS_NOT_NULL = 10, 11                                                                   # types.hpp:177-178
INVALID = 0xFFFFFFFF

_COMPARE = {EQUALS: np.equal, NOT_EQUALS: np.not_equal, LESS_THAN: np.less, LESS_THAN_EQUALS: np.less_equal,
            GREATER_THAN: np.greater, GREATER_THAN_EQUALS: np.greater_equal}


def decode(chunk, offsets=None):
    """(values, is_null) of a value or dictionary chunk, optionally at `offsets` (iterables' NULL conventions:
    value_column_iterable.hpp:90-100, dictionary_column_iterable.hpp:76-87)."""
    if chunk["kind"] == "value":
        v = chunk["values"]
        n = np.zeros(len(v), bool) if chunk.get("nulls") is None else chunk["nulls"].astype(bool)
    elif chunk["kind"] == "for":   # FrameOfReferenceColumn: block minimum + offset, 2048-row blocks (frame_of_reference_column.hpp:33-80)
        off = chunk["offsets"].astype(np.int64)
        v = (np.repeat(chunk["minima"].astype(np.int64), 2048)[:len(off)] + off).astype(chunk["minima"].dtype)
        n = np.zeros(len(v), bool) if chunk.get("nulls") is None else chunk["nulls"].astype(bool)
    elif chunk["kind"] == "rle":   # RunLengthColumn: run values + inclusive end positions (run_length_column.hpp:29-66)
        ends = chunk["end_positions"].astype(np.int64)
        lens = np.diff(np.concatenate([[-1], ends]))
        v = np.repeat(chunk["run_values"], lens)
        n = np.repeat(chunk["run_nulls"].astype(bool), lens) if chunk.get("run_nulls") is not None else np.zeros(len(v), bool)
    else:
        attr = chunk["attr"].astype(np.int64)
        n = attr == chunk["null_value_id"]
        d = chunk["dictionary"]
        if len(d):   # NULL decodes to T{} (dictionary_column_iterable.hpp:80): 0, or the empty string
            v = d[np.minimum(attr, len(d) - 1)].copy()
            v[n] = d.dtype.type()
        else:
            v = np.zeros(len(attr), np.int64)
    if offsets is not None:
        return v[offsets], n[offsets]
    return v, n


def cast_literal(value, dtype):
    """type_cast<ColumnDataType>(AllTypeVariant) (type_cast.hpp:39-58, called at single_column_table_scan_impl.cpp:36):
    floating column -> the literal as that floating type; integral column with a floating literal ->
    boost::numeric_cast<T>(double): truncation toward zero, bad_numeric_cast outside T's range."""
    dtype = np.dtype(dtype)
    if dtype.kind == "U":
        return str(value)   # lexical_cast<std::string>: an integer literal becomes its decimal text
    if dtype.kind == "f":
        return dtype.type(value)
    if isinstance(value, float):
        t = np.trunc(value)
        info = np.iinfo(dtype)
        if not (info.min <= t <= info.max):
            raise OverflowError("bad numeric conversion")
        return int(t)
    return int(value)


def _wide(values):
    """Integers are compared as int64 (no wrap-around on the literal), floats and strings in the column's own type
    (numpy orders str by code point = std::string's byte order on UTF-8)."""
    return values.astype(np.int64) if values.dtype.kind in "iu" else values


def _scan_value_like(values, is_null, predicate, value):
    """_unary_scan_with_value: NULL rows are skipped, the comparator decides the rest."""
    return np.flatnonzero(~is_null & _COMPARE[predicate](_wide(values), cast_literal(value, values.dtype)))


def _scan_dictionary(chunk, predicate, value):
    """single_column_table_scan_impl.cpp:83-199 on value ids."""
    d = _wide(chunk["dictionary"])
    attr = chunk["attr"].astype(np.int64)
    value = cast_literal(value, chunk["dictionary"].dtype)
    lower = int(np.searchsorted(d, value, side="left"))   # BaseDictionaryColumn::lower_bound
    upper = int(np.searchsorted(d, value, side="right"))  # upper_bound
    lower_id = INVALID if lower == len(d) else lower      # INVALID_VALUE_ID when no such entry
    upper_id = INVALID if upper == len(d) else upper
    search = lower_id if predicate in (EQUALS, NOT_EQUALS, LESS_THAN, GREATER_THAN_EQUALS) else upper_id  # :135-150
    unique = len(d)
    matches_all = {EQUALS: search != upper_id and unique == 1, NOT_EQUALS: search == upper_id,
                   LESS_THAN: search == INVALID, LESS_THAN_EQUALS: search == INVALID,
                   GREATER_THAN: search == 0, GREATER_THAN_EQUALS: search == 0}[predicate]             # :152-175
    matches_none = {EQUALS: search == upper_id, NOT_EQUALS: search == upper_id and unique == 1,
                    LESS_THAN: search == 0, LESS_THAN_EQUALS: search == 0,
                    GREATER_THAN: search == INVALID, GREATER_THAN_EQUALS: search == INVALID}[predicate]  # :177-199
    not_null = attr != chunk["null_value_id"]
    if matches_all:                                        # always_true over the attribute vector, NULLs still skipped
        return np.flatnonzero(not_null)
    if matches_none:
        return np.zeros(0, np.int64)
    # _with_operator_for_dict_column_scan (base_single_column_table_scan_impl.hpp): ==, !=, <, <, >=, >= on value ids
    op = {EQUALS: np.equal, NOT_EQUALS: np.not_equal, LESS_THAN: np.less, LESS_THAN_EQUALS: np.less,
          GREATER_THAN: np.greater_equal, GREATER_THAN_EQUALS: np.greater_equal}[predicate]
    return np.flatnonzero(not_null & op(attr, search))


def scan_chunk(chunk, predicate, value):
    """Ascending matching chunk offsets of one chunk. value None = NULL literal."""
    if predicate in (IS_NULL, IS_NOT_NULL):
        if chunk["kind"] == "reference":
            v, n = _dereference(chunk)
        else:
            v, n = decode(chunk)
        return np.flatnonzero(n if predicate == IS_NULL else ~n)
    if predicate not in _COMPARE:
        raise ValueError("Unsupported comparison type encountered")  # Fail(), :147
    if value is None:
        return np.zeros(0, np.int64)                       # :25-34
    if chunk["kind"] == "dict":
        return _scan_dictionary(chunk, predicate, value)
    if chunk["kind"] == "reference":
        v, n = _dereference(chunk)
        return _scan_value_like(v, n, predicate, value)
    v, n = decode(chunk)
    return _scan_value_like(v, n, predicate, value)


def _dereference(chunk):
    """(values, is_null) of a reference chunk: NULL RowID -> NULL (reference_column_iterable.hpp:59-61), else the
    referenced row's value / null flag."""
    pl = chunk["pos_list"]
    n_rows = len(pl)
    kinds = {c[{"value": "values", "dict": "dictionary", "for": "minima", "rle": "run_values"}[c["kind"]]].dtype
             for c in chunk["referenced"]}
    values = np.zeros(n_rows, kinds.pop() if len(kinds) == 1 else np.int64)
    is_null = np.zeros(n_rows, bool)
    for i in range(n_rows):
        c, o = int(pl[i, 0]), int(pl[i, 1])
        if c == INVALID and o == INVALID:
            is_null[i] = True
            continue
        v, n = decode(chunk["referenced"][c], np.array([o]))
        values[i], is_null[i] = v[0], bool(n[0])
    return values, is_null


def table_scan(chunks, chunk_ids, predicate, value, dereference=False):
    """The operator's result for one column: (pos (n, 2) uint32, chunk_begin (n_chunks + 1) uint32).
    dereference: reference chunks emit pos_list[offset] (table_scan.cpp:124-131) instead of (chunk_id, offset)."""
    pos, begin = [], [0]
    for chunk, cid in zip(chunks, chunk_ids):
        m = scan_chunk(chunk, predicate, value)
        if dereference and chunk["kind"] == "reference":
            rows = chunk["pos_list"][m].astype(np.uint32).reshape(-1, 2)
        else:
            rows = np.stack([np.full(len(m), cid, np.uint32), m.astype(np.uint32)], axis=1).reshape(-1, 2)
        pos.append(rows)
        begin.append(begin[-1] + len(m))
    return (np.concatenate(pos) if pos else np.zeros((0, 2), np.uint32)), np.array(begin, np.uint32)
