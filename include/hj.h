/*
 * hj.h — C ABI of libhj, the B200 (sm_100a) CUDA library behind Hyrise's JoinHash operator.
 *
 * This header is the whole drop-in boundary. The host shim that keeps Hyrise's
 * `JoinHash : AbstractJoinOperator` signature (src/lib/operators/join_hash.hpp:24-45 in the
 * reference) calls these functions from `_on_execute()`; nothing else of the reference is
 * rebuilt. Plain C: pointers and sizes only, no C++/torch types, no exceptions cross the ABI.
 *
 * What each entry point replaces in the reference (paths relative to /root/reference):
 *   hj_join          materialize_input        src/lib/operators/join_hash.cpp:203-285
 *                    partition_radix_parallel src/lib/operators/join_hash.cpp:287-355
 *                    build                    src/lib/operators/join_hash.cpp:127-185
 *                    probe                    src/lib/operators/join_hash.cpp:362-466
 *                    probe_semi_anti          src/lib/operators/join_hash.cpp:468-527
 *                    (orchestrated as JoinHashImpl::_on_execute, join_hash.cpp:686-800)
 *   hj_fetch         the per-partition PosLists handed to write_output_columns,
 *                    src/lib/operators/join_hash.cpp:829-855
 *   hj_translate     PosList dereferencing for reference-table inputs in write_output_columns,
 *                    src/lib/operators/join_hash.cpp:584-592
 *   hj_radix_bits    JoinHashImpl ctor heuristic, src/lib/operators/join_hash.cpp:654-668
 *                    (re-derived for 228 KB of shared memory instead of a 256 KB L2)
 *
 * There is NO CPU fallback in the library: every entry point that computes needs a CUDA device
 * and returns HJ_ERR_CUDA otherwise.
 */
#ifndef HJ_H_
#define HJ_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HJ_ABI_VERSION 3

/* ---- ABI constants mirrored from the reference ------------------------------------------- */

/* RowID, src/lib/types.hpp:97-99: {ChunkID uint32, ChunkOffset uint32}, 8 bytes.
 * PosList = contiguous vector<RowID> (types.hpp:138) so an hj_row_id* may alias PosList::data(). */
typedef struct hj_row_id {
  uint32_t chunk_id;
  uint32_t chunk_offset;
} hj_row_id;

#define HJ_INVALID_CHUNK_ID 0xFFFFFFFFu     /* types.hpp:95  INVALID_CHUNK_ID     */
#define HJ_INVALID_CHUNK_OFFSET 0xFFFFFFFFu /* types.hpp:94  INVALID_CHUNK_OFFSET */
/* NULL_ROW_ID = {INVALID_CHUNK_ID, INVALID_CHUNK_OFFSET}, types.hpp:150 */

/* JoinMode, src/lib/types.hpp:189 — same integer values. Outer/Cross are rejected
 * (join_hash.cpp never receives them: lqp_translator.cpp:329-336, abstract_join_operator.cpp:19). */
enum hj_join_mode {
  HJ_MODE_INNER = 0,
  HJ_MODE_LEFT = 1,
  HJ_MODE_RIGHT = 2,
  HJ_MODE_OUTER = 3,
  HJ_MODE_CROSS = 4,
  HJ_MODE_SEMI = 5,
  HJ_MODE_ANTI = 6
};

/* DataType of the key column (src/lib/all_type_variant.hpp:34-39). Int = int32, Long = int64, Float, Double.
 * The hashed type of a join follows JoinHashTraits (join_hash/hash_traits.hpp:14-35): two integer columns hash as the
 * wider integer; if either column is floating point the join hashes in the floating type (the wider one of two),
 * and integer keys are converted to it — the reference goes through boost::lexical_cast (type_cast.hpp:44-48),
 * which for an integer's exact decimal text is the correctly rounded conversion, i.e. the IEEE cast done here.
 * Float keys compare like in the reference: by value through the hash of their bytes, so NaN matches nothing and
 * -0.0 does not meet +0.0. A FLOAT x DOUBLE pair is widened with the IEEE cast, NOT with the reference's 9-digit
 * decimal round trip (the shim refuses that pair). Float keys: hj_join only (not hj_join_sharded). */
enum hj_key_type { HJ_KEY_INT32 = 0, HJ_KEY_INT64 = 1, HJ_KEY_FLOAT = 2, HJ_KEY_DOUBLE = 3 };

/* How one chunk's key column is encoded. */
enum hj_segment_kind {
  /* ValueColumn<T>, src/lib/storage/value_column.hpp:15-73: values T[n_rows] (+ 1 byte/row null flags) */
  HJ_SEG_VALUE = 0,
  /* DictionaryColumn<T>, src/lib/storage/dictionary_column.hpp:20-72 with a
   * FixedSizeByteAlignedVector<uint8|16|32> attribute vector
   * (storage/vector_compression/fixed_size_byte_aligned/fixed_size_byte_aligned_vector.hpp:21-51);
   * a row is NULL iff attr[i] == null_value_id (dictionary_column_iterable.hpp:76-80). */
  HJ_SEG_DICT = 1,
  /* ReferenceColumn, src/lib/storage/reference_column.hpp:19-50: the shim has gathered the keys
   * through the PosList into `values` (+ `nulls`); `row_offsets` optionally overrides the
   * ChunkOffset reported per row (quirk Q2, reference_column_iterable.hpp:60), NULL = 0,1,2,... */
  HJ_SEG_GATHERED = 2,
  /* ReferenceColumn handed over as it is stored (reference_column.hpp:19-50): `attr` is the PosList
   * (hj_row_id[n_rows], NULL_ROW_ID = NULL key), `values` points to an hj_relation describing the key
   * column of the REFERENCED data table (segs[c] = its chunk c, VALUE or DICT, same hj_location as this
   * relation). The library gathers (value, is_null) through the PosList on the device
   * (reference_column_iterable.hpp:59-90) — SURVEY §8(f) N1. Emitted RowIDs are (chunk_id, 0..n_rows-1),
   * to be mapped to the data table with hj_translate like those of GATHERED segments. */
  HJ_SEG_REFERENCE = 3,
  /* FrameOfReferenceColumn<int32|int64>, src/lib/storage/frame_of_reference_column.hpp:33-80 with a
   * FixedSizeByteAligned offset vector: `values` = block minima T[ceil(n_rows / 2048)], `attr` = offsets
   * uintW[n_rows] (attr_width 1/2/4), `nulls` = 1 byte per row or NULL (the shim expands the column's
   * vector<bool>). key = minima[i / 2048] + offsets[i] (frame_of_reference_iterable.hpp). SURVEY §8(f) N2. */
  HJ_SEG_FOR = 4,
  /* RunLengthColumn<T>, src/lib/storage/run_length_column.hpp:29-66: `values` = run values T[dict_n],
   * `attr` = inclusive end positions uint32[dict_n], `nulls` = 1 byte per RUN or NULL; dict_n = number of
   * runs. Row i belongs to the first run whose end position is >= i (run_length_column_iterable.hpp). */
  HJ_SEG_RLE = 5
};

/* Where the arrays a segment points to live. */
enum hj_location { HJ_LOC_HOST = 0, HJ_LOC_DEVICE = 1 };

typedef struct hj_segment {
  uint32_t chunk_id;      /* ChunkID reported in emitted RowIDs (join_hash.cpp:266)        */
  uint32_t n_rows;        /* rows in this chunk's column                                    */
  uint8_t kind;           /* hj_segment_kind                                                */
  uint8_t key_type;       /* hj_key_type of `values`                                        */
  uint8_t attr_width;     /* DICT/FOR: bytes per attribute-vector / offset entry: 1, 2 or 4
                             (FixedSizeByteAligned), or 0: `attr` is a SimdBp128 stream
                             (vector_compression/simd_bp128/simd_bp128_packing.hpp:17-21: per 2048
                             values one 16-byte word of 16 bit widths, then for each block of 128
                             values `width` 16-byte words, four interleaved 32-bit lanes) — N2  */
  uint8_t reserved0;
  uint32_t dict_n;        /* DICT: entries in the dictionary; RLE: number of runs           */
  uint32_t null_value_id; /* DICT: ValueID meaning NULL (== dict_n in the reference)        */
  uint32_t reserved1;
  const void* values;     /* VALUE/GATHERED: T[n_rows]; DICT: sorted dictionary T[dict_n];
                             REFERENCE: const hj_relation* of the referenced key column      */
  const uint8_t* nulls;   /* VALUE/GATHERED: 1 byte per row (non-zero = NULL) or NULL ptr   */
  const void* attr;       /* DICT: uintW[n_rows]; REFERENCE: hj_row_id[n_rows] (the PosList) */
  const uint32_t* row_offsets; /* GATHERED: ChunkOffset per row, or NULL for 0..n_rows-1   */
} hj_segment;

typedef struct hj_relation {
  const hj_segment* segs; /* one per chunk, in chunk order (host memory)                    */
  uint32_t n_segs;
  uint32_t location;      /* hj_location of the arrays the segments point to                */
  uint64_t n_rows;        /* sum of segs[i].n_rows                                          */
} hj_relation;

/* ---- options, results, timings ------------------------------------------------------------ */

/* Flags selecting between the reference's literal behaviour and SQL semantics for the quirks
 * SURVEY.md §8a lists. Default (0) reproduces the reference literally. */
#define HJ_FLAG_SQL_NULL_PROBE 0x1u /* Q4: a NULL probe key in Left/Right never matches; default:
                                       it is looked up as its stored value (VALUE/GATHERED) or
                                       T{} = 0 (DICT), join_hash.cpp:253,409 */
#define HJ_FLAG_TIMINGS 0x2u        /* record per-phase CUDA-event timings (adds event overhead) */
#define HJ_FLAG_XCHG_STAGED 0x4u    /* hj_join_sharded: exchange through send buffers + NCCL send/recv even when the
                                       ranks can map each other's memory (default: the scatter kernel stores straight
                                       into the peers' receive buffers over NVLink)                               */
#define HJ_FLAG_TWO_SWEEP_PARTITION 0x8u /* two-pass partitioning (radix_bits > 8) by the v2 kernels: histogram sweep
                                       that stores decoded keys, then two scatter passes. Default: pass 1 is done
                                       slab-locally inside the histogram kernel (hj_fused.cuh). Same results.    */

#define HJ_FLAG_XCHG_GENERAL 0x10u   /* hj_join_sharded: always the general exchange (count sweep, RowID-carrying tuples,
                                       hj_comm.cuh). Default: the push exchange (hj_xpush.cuh) wherever its preconditions
                                       hold, the general one otherwise. Same results.                            */

typedef struct hj_options {
  uint32_t abi_version;       /* HJ_ABI_VERSION                                              */
  uint32_t flags;             /* HJ_FLAG_*                                                   */
  int32_t radix_bits;         /* <0: library heuristic (hj_radix_bits); else forced (0..15).
                                 The reference ignores its ctor argument (quirk Q1).         */
  uint32_t reserved0;
  uint64_t output_capacity;   /* 0: estimate; else initial capacity of the result buffers in
                                 rows (the library grows and re-probes on overflow)          */
  uint64_t stream_slice_rows; /* streamed join (hj_bind_output): minimum probe rows per slice; 0: 32 Mi.
                                 hj_join_sharded: != 0 cuts every rank's probe shard into up to 8 slices of at
                                 least this many rows (same value on all ranks) that are exchanged one after the
                                 other, so the receive buffers hold two slices instead of the whole relation;
                                 0: one piece (faster) */
  uint64_t reserved[3];
} hj_options;

typedef struct hj_result_info {
  uint64_t n_rows;          /* emitted result rows (pairs, or single RowIDs for Semi/Anti)    */
  uint32_t n_sides;         /* 2 for Inner/Left/Right, 1 for Semi/Anti                        */
  uint32_t radix_bits;      /* bits actually used                                             */
  uint64_t n_build_tuples;  /* build rows that entered the partitioning (non-NULL keys)       */
  uint64_t n_probe_tuples;  /* probe rows that entered the partitioning                       */
  uint32_t n_partitions;
  uint32_t n_overflow_partitions; /* build partitions larger than one shared-memory table     */
  uint32_t n_probe_passes;  /* 1, or 2 when the first output capacity estimate overflowed     */
  uint32_t n_probe_slices;  /* > 0: the join was streamed in that many probe slices (hj_bind_output) */
} hj_result_info;

#define HJ_N_PHASES 10
/* phase indices of hj_timings.ms[] */
enum hj_phase {
  HJ_PH_H2D = 0,          /* staging host-resident segments                                 */
  HJ_PH_MATERIALIZE = 1,  /* decode + histogram, both relations                             */
  HJ_PH_SCATTER1 = 2,     /* radix pass 1, both relations                                   */
  HJ_PH_SCATTER2 = 3,     /* radix pass 2, both relations                                   */
  HJ_PH_BUILD_PROBE = 4,  /* shared-memory build + probe + emit                             */
  HJ_PH_FINALIZE = 5,     /* overflow-partition fix-up, NULL-probe emission                 */
  HJ_PH_D2H = 6,          /* hj_fetch copies                                                */
  HJ_PH_TOTAL = 7,        /* whole hj_join / hj_join_sharded, device side                   */
  HJ_PH_XCHG_PARTITION = 8, /* hj_join_sharded: split by destination rank (count + scatter)  */
  HJ_PH_XCHG_TRANSFER = 9   /* hj_join_sharded: all-to-all over NVLink                        */
};
typedef struct hj_timings {
  float ms[HJ_N_PHASES];
  uint64_t bytes_h2d;
  uint64_t bytes_d2h;
  uint32_t n_kernel_launches; /* kernels of this library launched by the last hj_join       */
  uint32_t xchg_path;         /* hj_join_sharded: 1 = general exchange (hj_comm.cuh), 2 = push exchange (hj_xpush.cuh); else 0 */
  uint64_t bytes_nvlink_sent; /* hj_join_sharded: bytes this rank sent to its peers         */
} hj_timings;

/* ---- status codes -------------------------------------------------------------------------- */
enum hj_status {
  HJ_OK = 0,
  HJ_ERR_INVALID = 1,     /* bad argument / unsupported mode or type (the shim turns it into Fail(), utils/assert.hpp:49) */
  HJ_ERR_CUDA = 2,        /* CUDA runtime error, no device                                   */
  HJ_ERR_OOM = 3,         /* device allocation failed                                        */
  HJ_ERR_CAPACITY = 4,    /* hj_fetch capacity smaller than the result                       */
  HJ_ERR_TOO_LARGE = 5,   /* a relation has >= 2^32-1 rows                                   */
  HJ_ERR_STATE = 6,       /* hj_fetch without a preceding successful hj_join                 */
  HJ_ERR_COMM = 7         /* multi-GPU exchange failed                                       */
};

typedef struct hj_ctx hj_ctx; /* opaque; all library state lives here (re-entrant, one ctx per calling thread) */

/* ---- entry points -------------------------------------------------------------------------- */

/* Library/ABI version of the loaded .so (== HJ_ABI_VERSION it was compiled with). */
int hj_abi_version(void);
/* sizeof() of the ABI structs as compiled into the library, for binding generators to self-check:
 * 0 hj_segment, 1 hj_relation, 2 hj_options, 3 hj_result_info, 4 hj_timings, 5 hj_row_id; 0 for other values. */
size_t hj_abi_sizeof(int which);

/* Creates a context on CUDA device `device_id`. `opts` may be NULL (defaults). */
int hj_create(int device_id, const hj_options* opts, hj_ctx** out_ctx);
void hj_destroy(hj_ctx* ctx);

/* Last error text of this context (never NULL). With ctx == NULL: error of the last failed hj_create
 * on this thread. */
const char* hj_last_error(const hj_ctx* ctx);

/* Radix bits the library would use for `n_build_rows` keys of `key_type`
 * (replaces join_hash.cpp:654-668). Pure host arithmetic. */
int hj_radix_bits(uint64_t n_build_rows, int key_type);

/* Runs the whole join on the device and keeps the result PosLists in device memory owned by ctx.
 * `build`/`probe` are the relations AFTER the operator's side selection (join_hash.cpp:58-76);
 * `mode` is the operator's JoinMode. NULL keys: dropped on the build side always; on the probe
 * side kept only for Left/Right (join_hash.cpp:715,755). Synchronous: returns when the result
 * count is known. `cuda_stream` is a cudaStream_t (NULL = the context's own stream). */
int hj_join(hj_ctx* ctx, const hj_relation* build, const hj_relation* probe, int mode,
            void* cuda_stream, hj_result_info* out_info);

/* Copies the result into caller-allocated HOST buffers (PosList::resize(n) then data()).
 * `build_side` receives the build relation's RowIDs (NULL_ROW_ID for unmatched probe rows in
 * Left/Right) and must be NULL for Semi/Anti; `probe_side` receives the probe relation's RowIDs.
 * `capacity` is in rows. */
int hj_fetch(hj_ctx* ctx, hj_row_id* build_side, hj_row_id* probe_side, uint64_t capacity);

/* Optional, before hj_join: binds caller-allocated HOST buffers (capacity in rows; PosList::reserve/resize to an
 * upper estimate, e.g. the probe row count of a foreign-key join) that the NEXT hj_join fills directly; the binding
 * then lapses. With host-resident inputs of many chunks the join is streamed: the probe relation is processed in
 * slices of whole chunks, and while one slice is joined the next one is copied in and the PosLists of the previous
 * one are copied out, so both PCIe directions and the GPU work at the same time (the operator's end-to-end time
 * is then that of the result download). Result rows of a slice are contiguous; order across slices follows the
 * probe chunks. After such a join hj_fetch with the same pointers is a no-op and hj_result_device has nothing to
 * return. HJ_ERR_CAPACITY if the result does not fit: join again unbound. `build_side` may be NULL for Semi/Anti.
 * Pinned (page-locked) buffers are needed for the copies to overlap. Replaces nothing in the reference — its
 * PosLists are written in place (join_hash.cpp:829-855); this is the device-side equivalent of that. */
int hj_bind_output(hj_ctx* ctx, hj_row_id* build_side, hj_row_id* probe_side, uint64_t capacity_rows);

/* Device pointers to the result of the last hj_join (valid until the next hj_join/hj_destroy).
 * `build_side` is set to NULL for Semi/Anti. */
int hj_result_device(hj_ctx* ctx, const hj_row_id** build_side, const hj_row_id** probe_side,
                     uint64_t* n_rows);

/* A7 for reference-table inputs (join_hash.cpp:584-592): inout[i] = pos_lists[inout[i].chunk_id]
 * [inout[i].chunk_offset], NULL_ROW_ID stays NULL_ROW_ID. All pointers are HOST memory; the gather
 * runs on the device. `pos_list_sizes[c]` bounds-checks chunk c. */
int hj_translate(hj_ctx* ctx, hj_row_id* inout, uint64_t n, const hj_row_id* const* pos_lists,
                 const uint32_t* pos_list_sizes, uint32_t n_chunks);

/* Per-phase timings of the last hj_join/hj_fetch (needs HJ_FLAG_TIMINGS for the per-phase split). */
int hj_get_timings(const hj_ctx* ctx, hj_timings* out);

/* Frees cached device scratch (the context stays usable). JoinHash::_on_cleanup, join_hash.cpp:89. */
int hj_release_scratch(hj_ctx* ctx);

/* ---- TableScan on the device --------------------------------------------- */
/* Replaces the single-column-vs-literal scan of the reference — TableScan::_on_execute
 * (src/lib/operators/table_scan.cpp:84-171) with SingleColumnTableScanImpl
 * (operators/table_scan/single_column_table_scan_impl.cpp:23-133) and IsNullTableScanImpl
 * (is_null_table_scan_impl.cpp) — for integer columns, so that a join input filtered by a scan never
 * leaves the GPU: the scan's PosLists stay in device memory and are handed to hj_join as
 * HJ_SEG_REFERENCE segments of a device-resident relation.
 * Values are the reference's PredicateCondition (src/lib/types.hpp:166-179). */
enum hj_predicate {
  HJ_PRED_EQUALS = 0,
  HJ_PRED_NOT_EQUALS = 1,
  HJ_PRED_LESS_THAN = 2,
  HJ_PRED_LESS_THAN_EQUALS = 3,
  HJ_PRED_GREATER_THAN = 4,
  HJ_PRED_GREATER_THAN_EQUALS = 5,
  HJ_PRED_BETWEEN = 6,  /* refused: the reference's single-column scan fails on it too (table_scan_test.cpp:236) */
  HJ_PRED_IN = 7,       /* refused */
  HJ_PRED_LIKE = 8,     /* refused (string columns) */
  HJ_PRED_NOT_LIKE = 9, /* refused */
  HJ_PRED_IS_NULL = 10,
  HJ_PRED_IS_NOT_NULL = 11
};

#define HJ_SCAN_SLOTS 4 /* results a context keeps at a time (e.g. the two filtered inputs of a join) */
#define HJ_SCAN_DEREFERENCE 0x1u /* hj_scan scan_flags: REFERENCE chunks emit pos_list[offset], not (chunk_id, offset) */
#define HJ_SCAN_VALUE_IS_DOUBLE 0x2u /* hj_scan scan_flags: `value` holds the bit pattern of a double literal (memcpy) */

typedef struct hj_scan_info {
  uint64_t n_rows;            /* matching rows over all chunks                                  */
  uint32_t n_chunks;          /* = column->n_segs                                               */
  uint32_t n_nonempty_chunks; /* chunks with at least one match (the reference drops the others) */
} hj_scan_info;

/* Scans one int32 / int64 / float / double column (VALUE / DICT / GATHERED / FOR / RLE chunks, host or device
 * resident; REFERENCE chunks) with `column <predicate> value`.
 * Result `slot` (0..HJ_SCAN_SLOTS-1) then holds, in device memory, one PosList per input chunk, back to back:
 * chunk c's matches are pos[chunk_begin[c] .. chunk_begin[c+1]) in ascending chunk offset, as
 * BaseTableScanImpl::_unary_scan appends them (for a reference-table input the reference groups the matches by
 * referenced chunk instead; its tests compare unordered). A match is RowID{segs[c].chunk_id, offset} — the row of the
 * INPUT table, which is what the operator needs to filter every distinct PosList of a reference-table input
 * (table_scan.cpp:112-131). With HJ_SCAN_DEREFERENCE a REFERENCE chunk emits its own PosList entry at that offset
 * instead, i.e. the output PosList itself when all columns share the scanned column's PosList (the usual case:
 * no multi-level referencing, table_scan.cpp:98-110).
 * NULL keys match nothing but IS NULL; `value_is_null` != 0 (comparison with a NULL literal) matches nothing at all
 * (single_column_table_scan_impl.cpp:25-34). A dictionary chunk is scanned on its value ids against the literal's
 * lower/upper bound in the chunk's dictionary (:83-133) — no per-row dictionary access.
 * The literal is an integer in `value`, or with HJ_SCAN_VALUE_IS_DOUBLE a double whose bits are in `value`; it is
 * converted to the column's type the way type_cast<ColumnDataType>() does (type_cast.hpp:39-58): to float / double
 * for a floating column (the chunks of such a column are compared as values, dictionary chunks included), and a
 * double literal on an integer column is truncated toward zero (boost::numeric_cast; `a < 457.9` scans `a < 457`).
 * HJ_ERR_INVALID: refused predicate, literal outside the integer column's range. */
int hj_scan(hj_ctx* ctx, const hj_relation* column, int predicate, int64_t value, int value_is_null, int slot,
            uint32_t scan_flags, void* cuda_stream, hj_scan_info* out_info);

/* The same scan for dictionary chunks of ANY value type (strings included), on their value ids alone: the caller
 * does what SingleColumnTableScanImpl does once per dictionary chunk (:135-150, BaseDictionaryColumn::lower_bound /
 * upper_bound — two binary searches in the chunk's sorted dictionary, on the host where the strings are) and passes
 * bounds[2c] = lower bound and bounds[2c+1] = upper bound value id of the literal in chunk c (the dictionary size or
 * INVALID_VALUE_ID 0xFFFFFFFF when there is no such entry); the device then compares the attribute vectors against
 * them exactly as hj_scan does (:83-133). Every chunk must be HJ_SEG_DICT with a FixedSizeByteAligned attribute
 * vector (attr_width 1/2/4) and its null_value_id; `values` / `dict_n` / `key_type` are ignored, the dictionaries
 * are never copied. bounds may be NULL for IS [NOT] NULL. Results are read like hj_scan's. scan_flags must be 0. */
int hj_scan_value_ids(hj_ctx* ctx, const hj_relation* column, int predicate, const uint32_t* bounds, int slot,
                      uint32_t scan_flags, void* cuda_stream, hj_scan_info* out_info);

/* Device pointers to the result in `slot` (valid until the next hj_scan into that slot, hj_scan_release or
 * hj_destroy): pos = hj_row_id[n_rows], chunk_begin = uint32[n_chunks + 1]. */
int hj_scan_result_device(hj_ctx* ctx, int slot, const hj_row_id** pos, const uint32_t** chunk_begin, uint64_t* n_rows,
                          uint32_t* n_chunks);

/* Copies the result in `slot` to host memory: pos[capacity_rows] and (optionally) chunk_begin[n_chunks + 1].
 * pos == NULL: only chunk_begin (the PosLists stay on the device for the operator that follows). */
int hj_scan_fetch(hj_ctx* ctx, int slot, hj_row_id* pos, uint64_t capacity_rows, uint32_t* chunk_begin);

/* Frees the result in `slot` (all slots if slot < 0). TableScan's output is owned by the operator, so it is NOT
 * touched by hj_release_scratch. */
int hj_scan_release(hj_ctx* ctx, int slot);

/* A7 for a result that stays in HBM (the device-side hj_translate): `device_inout` — n RowIDs in device memory,
 * e.g. a side of hj_result_device, MODIFIED IN PLACE — holds rows of `input`, the DEVICE-resident relation that
 * side was joined from. Every RowID whose chunk is an HJ_SEG_REFERENCE segment of `input` is replaced by that
 * segment's PosList entry, i.e. by the row of the referenced data table (join_hash.cpp:584-592); NULL_ROW_ID stays
 * NULL_ROW_ID, RowIDs of data chunks stay as they are. Afterwards the PosList addresses the data table and can be
 * handed to hj_gather_column for any of its columns. HJ_ERR_INVALID if a RowID points outside its chunk. */
int hj_translate_device(hj_ctx* ctx, hj_row_id* device_inout, uint64_t n, const hj_relation* input, void* cuda_stream);

/* ---- reading a column through a PosList on the device (consumers of the join) ---------------- */
/* The values of `column` (a DEVICE-resident data-table column: VALUE / DICT / FOR / RLE chunks, segs[i] = chunk i) at
 * the rows of a device-resident PosList — what every operator after the join (Aggregate, Projection, the next
 * join) does first when it reads a ReferenceColumn (reference_column_iterable.hpp:30-75: per RowID the referenced
 * chunk's value or NULL; NULL_ROW_ID = NULL). `device_pos` is typically a side of hj_result_device or a result of
 * hj_scan_result_device. Output, in device memory owned by ctx and valid until the next hj_gather_column into the
 * same `slot` (0..HJ_GATHER_SLOTS-1: e.g. a group-by column and an aggregate column side by side) /
 * hj_release_scratch / hj_destroy: out_values = n_rows values of the column's type (*out_key_type: int32 / int64 as
 * such, float / double as their IEEE bit patterns in 4 / 8 bytes, NaNs canonicalised; 0 where NULL) and out_nulls =
 * one byte per row. Runs the same kernel the join uses for REFERENCE key columns (k_gather_refs).
 * HJ_ERR_INVALID if a RowID points outside the column. */
#define HJ_GATHER_SLOTS 4
int hj_gather_column(hj_ctx* ctx, const hj_row_id* device_pos, uint64_t n_rows, const hj_relation* column, int slot,
                     void* cuda_stream, const void** out_values, const uint8_t** out_nulls, int* out_key_type);

/* Ungrouped aggregates of a column in device memory (typically hj_gather_column's output: the column of the joined
 * rows), all from one pass: what Aggregate computes per aggregate column when there are no GROUP BY columns
 * (operators/aggregate.cpp:203-252 with the functors :141-201 and result types aggregate_traits.hpp:16-66). */
typedef struct hj_aggregate {
  uint64_t n_rows;     /* COUNT(*)                                                                          */
  uint64_t n_values;   /* COUNT(column): rows that are not NULL                                             */
  int32_t is_floating; /* the column is float / double: read the *_double fields; else the *_int fields     */
  int32_t reserved;
  int64_t sum_int;     /* SUM of an int32 / int64 column, int64 arithmetic (wraps like the reference)       */
  double sum_double;   /* SUM of a float / double column in double; for integer columns (double)sum_int.    */
                       /* AVG = sum_double / n_values (aggregate.cpp:688-712)                               */
  int64_t min_int, max_int;       /* MIN / MAX of an integer column                                         */
  double min_double, max_double;  /* MIN / MAX of a float / double column (float values widened exactly)    */
} hj_aggregate;
/* device_values: n_rows values of type key_type (float / double as IEEE values or their bit patterns — the same
 * bytes); device_nulls: one byte per row or NULL. With n_values == 0 SUM / MIN / MAX / AVG are NULL in the
 * reference (:635-642) and the sum/min/max fields are 0. Double sums are folded in a fixed order (reproducible run
 * to run) that is not the reference's row order: they agree up to rounding. NaN values in MIN / MAX follow the
 * reference (aggregate.cpp:141-165: the first non-NULL value seeds the aggregate, later values replace it on plain
 * < / >): MIN and MAX are NaN exactly when the column's first non-NULL value is a NaN, otherwise NaNs are ignored;
 * the grouped entry points apply the same rule per group. */
int hj_aggregate_column(hj_ctx* ctx, const void* device_values, const uint8_t* device_nulls, uint64_t n_rows,
                        int key_type, void* cuda_stream, hj_aggregate* out);

/* hj_aggregate_column for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc). No counterpart in the reference.
 * Same computation and same kernels' arithmetic as hj_aggregate_column (A1 one partial per CTA, A2 fold in CTA order: the
 * results are identical, double sums included), but it calls no cudaMalloc / cudaFree / allocating routine, does not
 * synchronise and copies nothing to the host: the per-CTA partials live in a file-scope `static __device__` array of
 * libhj (HJ_STATIC_AGG_MAX_PARTIALS + 1 records of 64 bytes = 128 KiB, held by every GPU that loads the library until
 * the process exits), and the result is written to `device_out`, one hj_aggregate in DEVICE memory provided by the caller,
 * in stream order on `cuda_stream` (NULL = the context's own stream).
 * Limits: n_rows <= HJ_STATIC_AGG_MAX_ROWS, else HJ_ERR_TOO_LARGE with a message naming the limit; nothing is launched.
 * The scratch array is shared by every call on the device and keeps the last call's contents; nothing in it has to start
 * at zero (every partial that is read was written by the same call). Calls that overlap on the GPU (different streams,
 * different contexts or threads of one process) RACE on it: serialise them on one stream or with events. */
#define HJ_STATIC_AGG_MAX_PARTIALS 2048u
#define HJ_STATIC_AGG_MAX_ROWS (1ull << 31)
int hj_aggregate_column_static(hj_ctx* ctx, const void* device_values, const uint8_t* device_nulls, uint64_t n_rows,
                               int key_type, void* cuda_stream, hj_aggregate* device_out);

/* GROUP BY one integer column (Aggregate with one group-by column, aggregate.cpp:254-420): per distinct value of the
 * group column — NULL being a group of its own — the aggregates of hj_aggregate over the rows of that group.
 * device_group_keys / device_group_nulls: the group column of the n_rows input rows in device memory (int32 / int64;
 * typically hj_gather_column output in one slot), device_values / device_value_nulls / value_type: the aggregate
 * column in the same row order (another slot), or device_values == NULL for COUNT(*) alone. max_groups: an upper
 * bound of the number of groups if the caller has one (dictionary size, statistics): the hash table takes 64 bytes x 2 x
 * that, and HJ_ERR_CAPACITY is returned if there are more groups than promised. 0 = unknown: the table starts at 2^20
 * slots (64 MB; less for small inputs) and, whenever the pass finds it full, is rebuilt 8 x larger and the pass repeated
 * (at most 2 x n_rows slots, which always suffices). The groups stay on the device until
 * hj_aggregate_grouped_fetch copies them out (any order: the reference guarantees none, aggregate.hpp:45-47).
 * Double sums are accumulated with atomics: they agree with the reference up to rounding and are not bit-reproducible
 * from run to run. Several group-by columns: not built. */
typedef struct hj_group {
  int64_t key;         /* value of the group column (0 for the NULL group) */
  int32_t key_is_null;
  int32_t reserved;
  hj_aggregate agg;    /* agg.n_rows = COUNT(*) of the group */
} hj_group;
typedef struct hj_group_info {
  uint64_t n_groups;
  uint64_t n_rows;
} hj_group_info;
int hj_aggregate_grouped(hj_ctx* ctx, const void* device_group_keys, const uint8_t* device_group_nulls, int group_key_type,
                         const void* device_values, const uint8_t* device_value_nulls, int value_type, uint64_t n_rows,
                         uint64_t max_groups, void* cuda_stream, hj_group_info* out_info);
int hj_aggregate_grouped_fetch(hj_ctx* ctx, hj_group* groups, uint64_t capacity);
/* GROUP BY up to 4 columns of any numeric type: the groups are identified by a REPRESENTATIVE ROW, like the
 * reference's hash_entry.row_id (aggregate.cpp:222) — two rows are in one group iff all their group-by columns are
 * equal (raw value bits; NULL equals NULL). hj_group.key of the fetched groups is the index of one row of the group
 * (key_is_null = 0): the caller reads the group-by values of the output from its own columns at that row, which is
 * what _write_groupby_output does (:722-735). Everything else as hj_aggregate_grouped. */
typedef struct hj_group_column {
  const void* values;   /* device memory, n_rows values of key_type (float / double compared by their bits) */
  const uint8_t* nulls; /* device memory, one byte per row, or NULL                                         */
  int32_t key_type;
  int32_t reserved;
} hj_group_column;
int hj_aggregate_grouped_rows(hj_ctx* ctx, const hj_group_column* group_columns, uint32_t n_group_columns,
                              const void* device_values, const uint8_t* device_value_nulls, int value_type, uint64_t n_rows,
                              uint64_t max_groups, void* cuda_stream, hj_group_info* out_info);
/* ... and with the columns given as chunk descriptors of ONE table (same chunks, same rows per chunk; data or reference
 * table, host or device resident): the representative row of a group is a row number of the whole table (chunks in
 * order). value_column == NULL: COUNT(*) per group. Uses the buffers of gather slot 0. */
int hj_aggregate_grouped_rows_relation(hj_ctx* ctx, const hj_relation* const* group_columns, uint32_t n_group_columns,
                                       const hj_relation* value_column, uint64_t max_groups, void* cuda_stream,
                                       hj_group_info* out_info);
/* The same with both columns given as chunk descriptors of ONE table (same chunks, same rows per chunk; data or
 * reference table, host or device resident) — the call the host-side Aggregate operator makes per aggregate column.
 * value_column == NULL: COUNT(*) per group. Uses the buffers of gather slots 0 and 1. */
int hj_aggregate_grouped_relation(hj_ctx* ctx, const hj_relation* group_column, const hj_relation* value_column,
                                  uint64_t max_groups, void* cuda_stream, hj_group_info* out_info);

/* The same aggregates over a column given as chunk descriptors — the input format of hj_join and hj_scan: VALUE /
 * DICT / FOR / RLE chunks of a data table or REFERENCE chunks of a reference table (e.g. a join's or scan's output
 * table), host or device resident — without materialising the column: what Aggregate::_aggregate_column does over
 * all chunks of one aggregate column (aggregate.cpp:203-252). This is the call the host-side Aggregate operator
 * makes for a plan `Aggregate(JoinHash(...))` or `Aggregate(TableScan(...))` without GROUP BY. */
int hj_aggregate_relation(hj_ctx* ctx, const hj_relation* column, void* cuda_stream, hj_aggregate* out);

/* ---- multi-GPU (one process per GPU) ---------------------------------------- */
/* The exchange has no reference counterpart: both relations are split by a hash of the key so that
 * every key's rows meet on one rank, exchanged over NVLink by the library's own kernels (stores into
 * peer-mapped memory; NCCL carries the control words, or — HJ_FLAG_XCHG_STAGED — the data), and
 * joined locally. RowIDs stay global, so per-rank results simply concatenate. */

/* The rank (0..n_ranks-1) on which all rows with this key meet. Pure host arithmetic; the same function runs on the
 * device. `key_type` is the HASHED key type of the join (HJ_KEY_INT64 if either column is int64). */
uint32_t hj_exchange_rank_of(int64_t key, int key_type, uint32_t n_ranks);

#define HJ_NCCL_UNIQUE_ID_BYTES 128
/* Rank 0 calls this and broadcasts the 128 bytes to the other ranks out of band. */
int hj_comm_unique_id(uint8_t id[HJ_NCCL_UNIQUE_ID_BYTES]);
/* Collective over all ranks: attaches an NCCL communicator to ctx. */
int hj_comm_init(hj_ctx* ctx, const uint8_t id[HJ_NCCL_UNIQUE_ID_BYTES], int n_ranks, int rank);
/* Collective: like hj_join, but `build`/`probe` are this rank's shard of the relations (segments
 * carry GLOBAL chunk ids); afterwards hj_fetch/hj_result_device return this rank's share of the
 * result. out_info counts are per rank. Two exchanges, same results (hj_timings.xchg_path): the PUSH
 * exchange for large joins over regular shards (VALUE / DICT chunks of 2^k rows with consecutive chunk
 * ids; one sweep, finished tuples stored into the peers' memory, no count pass) and the GENERAL one
 * for everything else. hj_bind_output before the call works as for hj_join; when every rank binds host
 * outputs for host-resident shards of a large join, the join is streamed in probe slices (copy-in,
 * exchange + join and copy-out overlap; the same number of slices on every rank). A rank-local failure
 * before the exchange (allocation, IPC mapping) is agreed on by all ranks: all of them return an error
 * (the failing rank its own, the others HJ_ERR_COMM) instead of one leaving the others in a collective. */
int hj_join_sharded(hj_ctx* ctx, const hj_relation* build_shard, const hj_relation* probe_shard,
                    int mode, void* cuda_stream, hj_result_info* out_info);

#ifdef __cplusplus
}
#endif
#endif /* HJ_H_ */
