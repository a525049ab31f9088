/*
 * arrow1_b200.h — C ABI of the B200-native (sm_100a) Filter / SortIndices / grouped-aggregate
 * hot path of Arrow C++ 5.0.0-SNAPSHOT (reference: zhztheplayer/arrow-1).
 *
 * This is the drop-in boundary: plain pointers and sizes, no Arrow and no torch types.  Every
 * pointer inside a b200_col, and every `out_*` / `scratch` pointer, is a DEVICE pointer in the
 * address space of the context's device unless the parameter is documented as host.  The
 * reference interface each entry point replaces is cited as file:line under
 * /root/reference/cpp/src/arrow/ .
 *
 * Conventions
 *  - every function returns a b200_status; the values equal arrow::StatusCode (status.h:78-97)
 *    so the Arrow-side glue maps them with a cast.  b200_last_error() holds the message.
 *  - all work is enqueued on the context's CUDA stream.  A call returns without synchronising
 *    unless it hands a host-visible scalar back (an output length, num_groups), in which case it
 *    synchronises that stream only.
 *  - bitmaps are Arrow bitmaps: LSB-first, bit i of the array is bit (offset + i) of the buffer.
 *  - no CPU fallback exists: without a CUDA device every compute entry point fails.
 */
#ifndef ARROW1_B200_H
#define ARROW1_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200_API __attribute__((visibility("default")))
#else
#define B200_API
#endif

/* == arrow::StatusCode, status.h:78-97 */
typedef enum b200_status {
  B200_OK = 0,
  B200_OUT_OF_MEMORY = 1,
  B200_KEY_ERROR = 2,
  B200_TYPE_ERROR = 3,
  B200_INVALID = 4,
  B200_IO_ERROR = 5,
  B200_CAPACITY_ERROR = 6,
  B200_INDEX_ERROR = 7,
  B200_UNKNOWN_ERROR = 9,
  B200_NOT_IMPLEMENTED = 10
} b200_status;

/* == arrow::Type::type for the fixed-width primitives on the path, type_fwd.h (enum Type) */
typedef enum b200_type {
  B200_BOOL = 1,
  B200_UINT8 = 2,
  B200_INT8 = 3,
  B200_UINT16 = 4,
  B200_INT16 = 5,
  B200_UINT32 = 6,
  B200_INT32 = 7,
  B200_UINT64 = 8,
  B200_INT64 = 9,
  B200_FLOAT = 11,
  B200_DOUBLE = 12,
  /* binary / utf8 with int32 offsets (BinaryType, StringType).  GROUP KEYS ONLY (VarLengthKeyEncoder,
   * hash_aggregate.cc:236-311): `data` = the int32 offsets buffer (length + 1 entries past `offset`),
   * `extra` = the value bytes. */
  B200_BINARY = 13
} b200_type;

/* Device column descriptor: mirror of compute::internal::PrimitiveArg
 * (compute/kernels/util_internal.h:35-47) over ArrayData (array/data.h:73-245), EXCEPT that
 * `data` is the start of buffers[1] and is NOT pre-shifted by `offset` (simpler for the glue:
 * data = buffers[1]->address(), offset = ArrayData::offset for every bit width). */
typedef struct b200_col {
  const void* validity; /* device; NULL = no validity bitmap (== no nulls)                 */
  const void* data;     /* device; values buffer (bitmap when type == B200_BOOL)          */
  int32_t type;         /* b200_type                                                       */
  int32_t reserved;
  int64_t length;       /* elements                                                        */
  int64_t offset;       /* elements (bits for bitmaps)                                     */
  int64_t null_count;   /* -1 = unknown (kUnknownNullCount); ignored if validity == NULL   */
  const void* extra;    /* device; B200_BINARY: the value bytes (buffers[2]); NULL otherwise */
} b200_col;

/* Mutable twin used for outputs the library fills. */
typedef struct b200_out_col {
  void* validity; /* device; may be NULL where the operation documents it */
  void* data;     /* device */
  int32_t type;
  int32_t reserved;
  int64_t length;
  int64_t null_count; /* filled by the library: exact count or -1 */
  void* extra;        /* device; B200_BINARY outputs: the value bytes (`data` = int32 offsets, length + 1 entries) */
} b200_out_col;

typedef struct b200_ctx b200_ctx;

/* ---- context ------------------------------------------------------------------------------
 * One context = one device + one stream (SURVEY §8b "Threading").  `cu_stream` is a
 * cudaStream_t / CUstream owned by the caller (NULL = the legacy default stream).
 * Replaces: the role ExecContext plays for CPU kernels (compute/exec.h:58-120) plus the
 * CudaContext a CudaBuffer belongs to (gpu/cuda_context.h:211-308). */
B200_API int b200_ctx_create(int device, void* cu_stream, b200_ctx** out);
B200_API int b200_ctx_destroy(b200_ctx* ctx);
B200_API int b200_ctx_set_stream(b200_ctx* ctx, void* cu_stream);
/* Waits for the context's stream.  Also the point where a kernel that had to give up is reported: the sort's
 * decoupled look-back bails out instead of spinning forever when a preceding tile never publishes (a preempted or
 * time-sliced GPU) and flags the context; synchronize then returns B200_UNKNOWN_ERROR and clears the flag.  Callers
 * that consume results through their own stream synchronisation should call this once per batch of work. */
B200_API int b200_ctx_synchronize(b200_ctx* ctx);
/* Execution knobs of a context (no reference counterpart; results never depend on them).  Names:
 *   "sort_high_word_first"   b200_sort_indices of a 64-bit column may sort by the upper key word alone (4 passes) and
 *                            repair the ties afterwards: 0 = decided per call from the digit histograms (default),
 *                            1 = always try, -1 = never
 * B200_KEY_ERROR for an unknown name. */
B200_API int b200_ctx_set_option(b200_ctx* ctx, const char* name, int64_t value);
/* test hook: sets the flag b200_ctx_synchronize reports */
B200_API int b200_ctx_inject_async_error(b200_ctx* ctx, uint32_t code);
/* message of the last failure on this context ("" if none); valid until the next call */
B200_API const char* b200_last_error(b200_ctx* ctx);
/* number of kernels this context has launched since creation (bench.py's gpu_launches) */
B200_API int64_t b200_ctx_launch_count(b200_ctx* ctx);
B200_API const char* b200_version(void);
/* Optional measurement hook: when enabled every hot kernel launch is bracketed by CUDA events on
 * the context's stream; b200_ctx_kernel_time sums the device time of the launches of one kernel
 * ("filter_plan", "filter_apply", "sort_hist", "sort_pass", "groupby_consume") since timing was
 * (re)enabled.  Synchronises. */
B200_API int b200_ctx_set_kernel_timing(b200_ctx* ctx, int enabled);
B200_API int b200_ctx_kernel_time(b200_ctx* ctx, const char* name, double* total_ms, int64_t* launches);

/* ---- Filter -------------------------------------------------------------------------------
 * null_selection: 0 = FilterOptions::DROP, 1 = FilterOptions::EMIT_NULL (api_vector.h:37-49).
 *
 * b200_filter_count     == GetFilterOutputSize                 (kernels/vector_selection.cc:61-88)
 * b200_filter_plan      == the same scan, keeping per-tile output offsets on the device so that
 *                          several columns can be compacted by one mask (FilterRecordBatch /
 *                          FilterTable, vector_selection.cc:1776-1866, reuse one index vector)
 * b200_filter_apply     == PrimitiveFilterImpl<T>::Exec        (vector_selection.cc:570-764)
 * b200_filter_out_spec  == the null_count / allocate_validity rules of PrimitiveFilter
 *                                                              (vector_selection.cc:799-812)
 * Caller flow (== PrimitiveFilter, vector_selection.cc:786-835): plan -> allocate -> apply.
 */
B200_API int b200_filter_plan_bytes(int64_t mask_length, size_t* out_bytes);
B200_API int b200_filter_count(b200_ctx* ctx, const b200_col* mask, int null_selection,
                               int64_t* out_len /* host */);
/* `plan` is device scratch of b200_filter_plan_bytes(mask->length) bytes.  If out_len (host) is
 * non-NULL the stream is synchronised and the output length returned; if NULL the call is
 * fully asynchronous (the length stays on the device inside the plan). */
B200_API int b200_filter_plan(b200_ctx* ctx, const b200_col* mask, int null_selection, void* plan,
                              int64_t* out_len /* host, nullable */);
/* Compacts `values` by the planned mask.  out_capacity = number of elements the output buffers
 * hold: the planned out_len, or mask->length when the caller skipped the host read-back.
 * out_data holds out_capacity elements of the value width (a bitmap for B200_BOOL);
 * out_validity holds out_capacity bits, or is NULL when b200_filter_out_spec says no validity
 * is allocated.  Output buffers are 8-byte aligned and bitmap buffers are sized to a multiple of
 * 8 bytes (Arrow pads to 64); bitmaps come back zero-padded, like KernelContext::AllocateBitmap
 * (compute/kernel.cc:51-59). */
B200_API int b200_filter_apply(b200_ctx* ctx, const b200_col* values, const b200_col* mask,
                               int null_selection, const void* plan, void* out_data,
                               void* out_validity /* nullable */, int64_t out_capacity);
B200_API int b200_filter_out_spec(const b200_col* values, const b200_col* mask, int null_selection,
                                  int* allocate_validity, int64_t* out_null_count);
/* Filter without a host round trip, for callers that already own an output of `out_capacity` rows (a reused buffer
 * sized by the mask length or by the previous batch): b200_filter_plan and b200_filter_apply back to back on the
 * stream.  == GetFilterOutputSize + PrimitiveFilterImpl::Exec (vector_selection.cc:61-88, 570-764) without
 * PreallocateData in between.  `scratch`: DEVICE, b200_filter_fused_scratch_bytes(mask length) bytes (a plan); its
 * first 8 bytes receive the int64 output length (b200_filter_plan_length_ptr).  out_len: HOST, nullable; when given
 * the call synchronises and returns the length.  If the selection emits more than out_capacity rows nothing is
 * written and the context reports it at the next b200_ctx_synchronize. */
B200_API int b200_filter_fused_scratch_bytes(int64_t mask_length, size_t* out_bytes);
B200_API int b200_filter_fused(b200_ctx* ctx, const b200_col* values, const b200_col* mask, int null_selection,
                               void* scratch, void* out_data, void* out_validity /* nullable */,
                               int64_t out_capacity, int64_t* out_len /* host, nullable */);
/* b200_filter_fused for callers that must not allocate and own no scratch (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, results and error behaviour as b200_filter_fused minus `scratch`: the plan lives
 * in a static device array inside the library (272 KB on every GPU that loads the library, held until the process
 * exits), and the call makes no cudaMalloc / cudaFree / allocating library call.
 * Limit: mask->length <= b200_filter_fused_noalloc_max_rows() (2^28 = 268 435 456 rows); longer masks fail with
 * B200_CAPACITY_ERROR and nothing is launched.
 * All calls on a device share the one array: calls that may overlap on the GPU (two contexts / streams on the same
 * device) race on it and must be ordered by the caller; calls on one context are ordered by its stream.  The array
 * keeps the previous call's contents; every call re-initialises what it reads. */
B200_API int64_t b200_filter_fused_noalloc_max_rows(void);
B200_API int b200_filter_fused_noalloc(b200_ctx* ctx, const b200_col* values, const b200_col* mask,
                                       int null_selection, void* out_data, void* out_validity /* nullable */,
                                       int64_t out_capacity, int64_t* out_len /* host, nullable */);
/* device address of the int64 output length inside a plan (for fully asynchronous callers) */
B200_API const int64_t* b200_filter_plan_length_ptr(const void* plan);

/* ---- mask producer -------------------------------------------------------
 * == the array-vs-scalar form of the comparison functions "equal", "not_equal", "greater",
 * "greater_equal", "less", "less_equal" (compute/kernels/scalar_compare.cc; op = 0..5 in that order) on a
 * numeric column: out bit i = values[i] OP *scalar, with the C++ operators of the physical type (a NaN
 * compares false except under not_equal).  `scalar` is a HOST pointer to one value of the column's type.
 * out_bitmap: device, ceil(length / 32) whole 32-bit words, bit offset 0.  Null-ness is not touched: like
 * the reference's null propagation (exec.cc PropagateNulls) the result's validity is the input's, and the
 * binding shares that buffer.  This is what feeds Filter in the Q1-shaped pipeline (config C5). */
B200_API int b200_compare_scalar(b200_ctx* ctx, const b200_col* values, int op,
                                 const void* scalar /* host */, void* out_bitmap);

/* ---- derived-column producers ------------------------------------------------
 * == "add", "subtract", "multiply" of compute/kernels/scalar_arithmetic.cc (the unchecked forms: integers
 * wrap like the reference's unsigned-cast arithmetic, :75-121, floats are plain IEEE operations, never
 * contracted into an FMA; a NaN result carries the GPU's canonical NaN payload) on numeric columns of ONE type.  Either side may be a scalar (a HOST pointer to one
 * value of that type, with that side's column pointer NULL); at least one side is a column.
 *   op: 0 add, 1 subtract, 2 multiply.
 *   a / b: column, or NULL when that operand is the scalar a_scalar / b_scalar.
 *   out_data: device, length * width bytes, every slot computed (also under nulls, like ScalarBinary).
 *   out_validity: device bitmap at bit offset 0, ceil(length / 32) whole words — the AND of the operands'
 *     validity (NullHandling::INTERSECTION, exec.cc PropagateNulls); pass NULL when neither operand has
 *     nulls.  The result's null_count is left to the caller (-1 = unknown, as the reference leaves it). */
typedef enum b200_arith_op { B200_ADD = 0, B200_SUBTRACT = 1, B200_MULTIPLY = 2 } b200_arith_op;
B200_API int b200_arithmetic(b200_ctx* ctx, int op, const b200_col* a, const void* a_scalar /* host */,
                             const b200_col* b, const void* b_scalar /* host */, void* out_data,
                             void* out_validity /* nullable */);

/* ---- SortIndices --------------------------------------------------------------------------
 * order: 0 = SortOrder::Ascending, 1 = SortOrder::Descending (api_vector.h:61-66).
 * == ArraySortIndices<UInt64Type, T>::Exec with ArrayCompareSorter / ArrayCountOrCompareSorter
 *    (kernels/vector_sort.cc:374-575): a STABLE sort of the row numbers; non-null non-NaN values
 *    first, then NaNs, then nulls, the last two in row order, for both orders.
 * out_indices: uint64[values->length] (device), never null (OUTPUT_NOT_NULL, vector_sort.cc:1747).
 * `index_base` is added to every emitted index (ChunkedArraySorter's `offset`, vector_sort.cc:657-832;
 * also the shard base of the multi-GPU sample sort).  scratch: device, >= *_scratch_bytes. */
B200_API int b200_sort_indices_scratch_bytes(int64_t length, size_t* out_bytes);
B200_API int b200_sort_indices(b200_ctx* ctx, const b200_col* values, int order,
                               uint64_t index_base, uint64_t* out_indices, void* scratch,
                               size_t scratch_bytes);
/* Which plan the last b200_sort_indices call that used `scratch` took (no reference counterpart; for benchmarks and
 * tests): *out_high_word_first = the passes ran over the upper key word only and ties were repaired afterwards;
 * *out_fell_back = that attempt met a tie run beyond the repair's limit and the ordinary 8-pass plan ran after it.
 * Synchronises the context. */
B200_API int b200_sort_indices_plan(b200_ctx* ctx, const void* scratch, int* out_high_word_first, int* out_fell_back);
/* ChunkedArray input (ChunkedArraySorter, vector_sort.cc:657-832): the stable sort of the LOGICAL
 * CONCATENATION of the chunks — indices are global, nulls / NaNs of every chunk come last.  Each chunk's
 * radix keys and global row numbers are written side by side, pair-sorted by key, and one more stable pass
 * on a 2-bit class key separates genuine maxima from NaNs from nulls.  out_indices: uint64[sum of lengths]. */
B200_API int b200_sort_indices_chunked_scratch_bytes(int64_t total_length, size_t* out_bytes);
B200_API int b200_sort_indices_chunked(b200_ctx* ctx, const b200_col* chunks, int n_chunks, int order,
                                       uint64_t* out_indices, void* scratch, size_t scratch_bytes);
/* Several sort keys: == SortIndices(RecordBatch | Table, SortOptions{sort_keys}) (SortIndicesMetaFunction,
 * vector_sort.cc:1632-1711; RadixRecordBatchSorter :955-1048, MultipleKeyRecordBatchSorter :1184-1356,
 * MultipleKeyComparator :1050-1181).  keys[0] is the most significant key; orders[k] as in b200_sort_indices.
 * Rows are ordered by key 0, ties by key 1, ...; within every key values come first (by its order), then NaNs,
 * then nulls; rows equal under all keys stay in row order (the reference's sorters are stable).  Done as an LSD
 * over the keys — one stable radix sort per key, least significant first — instead of one comparison sort with a
 * k-column comparator.  All keys have the same length (< 2^30 rows); n_keys == 1 is b200_sort_indices. */
B200_API int b200_sort_indices_multi_scratch_bytes(int64_t length, size_t* out_bytes);
B200_API int b200_sort_indices_multi(b200_ctx* ctx, const b200_col* keys, const int* orders, int n_keys,
                                     uint64_t* out_indices, void* scratch, size_t scratch_bytes);
/* The same over chunked key columns (SortIndices(Table), MultipleKeyTableSorter, vector_sort.cc:1381-1560):
 * keys[k * n_chunks + c] is chunk c of sort key k; chunk c has the same length under every key (the binding
 * slices the columns to common boundaries).  Indices address the logical concatenation. */
B200_API int b200_sort_indices_multi_chunked_scratch_bytes(int64_t total_length, size_t* out_bytes);
B200_API int b200_sort_indices_multi_chunked(b200_ctx* ctx, const b200_col* keys, const int* orders, int n_keys,
                                             int n_chunks, uint64_t* out_indices, void* scratch,
                                             size_t scratch_bytes);
/* Sample-sort building blocks for the multi-GPU path.  sort_keys writes the
 * order-preserving 64-bit radix key of every row (the same transform the local sort uses;
 * NaN -> 2^64-2, null -> 2^64-1) so that ranks can exchange (key, global row) pairs;
 * sort_pairs stably sorts such pairs by key, ties by their input order. */
B200_API int b200_sort_keys(b200_ctx* ctx, const b200_col* values, int order, uint64_t* out_keys);
/* As b200_sort_keys, and also writes out_rows[i] = (index_base + i) | class << 62 with class
 * 0 = value, 1 = NaN, 2 = null: the (key, global row) pairs a rank sends in the sample sort, the
 * class letting the receiver tell a null from a genuine maximum key. */
B200_API int b200_sort_keys_rows(b200_ctx* ctx, const b200_col* values, int order, uint64_t index_base,
                                 uint64_t* out_keys, uint64_t* out_rows /* nullable */);
/* Stable range partition for the sample sort: dest(row) = number of splitters <= key(row), with the keys /
 * rows of b200_sort_keys_rows (so NaN and null keys go to the last destination when the splitters are
 * clamped to 2^64-2).  Destination q's (key, row) pairs are written contiguously from sum(out_counts[0..q))
 * IN ROW ORDER, which is what lets the receiving rank finish with one stable b200_sort_pairs.
 * splitters: HOST, n_parts-1 ascending radix keys; n_parts <= 16; out_counts: HOST int64[n_parts]. */
B200_API int b200_sort_partition_scratch_bytes(int64_t length, size_t* out_bytes);
B200_API int b200_sort_range_partition(b200_ctx* ctx, const b200_col* values, int order,
                                       uint64_t index_base, const uint64_t* splitters /* host */,
                                       int n_parts, uint64_t* out_keys, uint64_t* out_rows,
                                       int64_t* out_counts /* host */, void* scratch, size_t scratch_bytes);
B200_API int b200_sort_pairs_scratch_bytes(int64_t length, size_t* out_bytes);
B200_API int b200_sort_pairs(b200_ctx* ctx, const uint64_t* keys, const uint64_t* payload,
                             int64_t length, uint64_t* out_keys /* nullable */,
                             uint64_t* out_payload, void* scratch, size_t scratch_bytes);
/* As b200_sort_pairs with a 32-bit payload that travels through the passes (no gather at the end):
 * out_payload[i] = (uint64_t)payload32[perm[i]].  Scratch: b200_sort_pairs_scratch_bytes. */
/* b200_sort_pairs for keys the caller knows to be below 2^key_bits (e.g. row numbers): only the digits that can
 * vary are histogrammed and sorted, and with key_bits <= 32 every pass carries 32-bit keys.  No sorted keys out. */
B200_API int b200_sort_pairs_bits(b200_ctx* ctx, const uint64_t* keys, const uint64_t* payload /* nullable */,
                                  int64_t length, int key_bits, uint64_t* out_payload, void* scratch,
                                  size_t scratch_bytes);
B200_API int b200_sort_pairs32(b200_ctx* ctx, const uint64_t* keys, const uint32_t* payload32,
                               int64_t length, uint64_t* out_keys /* nullable */,
                               uint64_t* out_payload, void* scratch, size_t scratch_bytes);

/* ---- Take (the gather that follows SortIndices) -----------------------------------
 * == compute::Take(values, indices, TakeOptions(boundscheck)) on a fixed-width column: PrimitiveTake /
 * PrimitiveTakeImpl / BooleanTakeImpl (kernels/vector_selection.cc:273-469, 493-526).
 *   out[i] = values[indices[i]]; a null index or a null value gives a null whose data is zero.
 *   indices: any integer type (the uint64 output of b200_sort_indices included), optional validity.
 *   out_data: indices->length elements of the value width (ceil(length / 32) whole words for B200_BOOL).
 *   out_validity: REQUIRED, ceil(length / 32) whole 32-bit words: like the reference, the bitmap is always
 *     produced.
 *   boundscheck != 0: a valid index that is negative or >= values->length fails with B200_INDEX_ERROR
 *     "Index N out of bounds" (CheckIndexBounds, util/int_util.cc) — this reads a flag back and synchronises.
 *     boundscheck == 0 matches TakeOptions::NoBoundsCheck(): such an index is undefined behaviour upstream;
 *     here the row comes out null.
 *   out_null_count (host, nullable): exact null count of the result (synchronises). */
B200_API int b200_take(b200_ctx* ctx, const b200_col* values, const b200_col* indices, int boundscheck,
                       void* out_data, void* out_validity, int64_t* out_null_count /* host, nullable */);
/* ChunkedArray values (TakeCA / TakeCC / TakeTA / TakeTC, vector_selection.cc:1921-2017): indices address the
 * LOGICAL CONCATENATION of the chunks.  Upstream concatenates the chunks on the host and calls array_take; here
 * each index is mapped to (chunk, row) by a search over the chunk boundaries and gathered from where the chunk
 * lies — no concatenated copy of the column is made.  All chunks share one type; everything else as b200_take. */
B200_API int b200_take_chunked(b200_ctx* ctx, const b200_col* chunks, int n_chunks, const b200_col* indices,
                               int boundscheck, void* out_data, void* out_validity,
                               int64_t* out_null_count /* host, nullable */);

/* ---- Range partition fused with the exchange (multi-GPU sample sort over NVLink peer memory) ----
 * The partition's scatter kernel stores every destination's (key, row) run straight into THAT rank's receive
 * buffer through a peer mapping, so the all-to-all is the scatter itself: no staging copy, no separate
 * collective pass.  Two phases, like the sizing calls elsewhere in this header:
 *   count : per-destination row counts of this rank's column (HOST int64[n_parts]); ranks exchange them,
 *           each rank derives where its run starts inside every destination buffer (sum of the counts of
 *           lower ranks for that destination);
 *   push  : the stable scatter, destination q's run written at dst_keys[q] + dst_offsets[q] (and the same
 *           element offset in dst_rows[q]).  rows32 != 0 stores rows as uint32 = (index_base + i) | class << 30
 *           (requires index_base + length <= 2^30), else uint64 with the class at bit 62.
 * push must follow count with the same column, order, splitters and scratch.  The caller orders the
 * exchange: no rank may push before every destination has finished reading its buffer's previous
 * contents, and no rank may read its buffer before every rank's push has completed (stream-ordered
 * collectives on either side do both; see arrow1_b200/dist.py). */
B200_API int b200_sort_range_partition_count(b200_ctx* ctx, const b200_col* values, int order,
                                             const uint64_t* splitters /* host */, int n_parts,
                                             int64_t* out_counts /* host */, void* scratch, size_t scratch_bytes);
B200_API int b200_sort_range_partition_push(b200_ctx* ctx, const b200_col* values, int order,
                                            uint64_t index_base, const uint64_t* splitters /* host */, int n_parts,
                                            uint64_t* const* dst_keys /* host array of device pointers */,
                                            void* const* dst_rows /* host array of device pointers */,
                                            const int64_t* dst_offsets /* host, elements */, int rows32,
                                            void* scratch, size_t scratch_bytes);
/* Exchange buffers other processes of this node can map: plain device allocations exported as CUDA IPC
 * handles (64 opaque bytes, exchanged by the caller, e.g. with an all-gather) and opened by the peers.
 * The one place this library allocates device memory, because an IPC handle names a whole allocation. */
#define B200_PEER_HANDLE_BYTES 64
B200_API int b200_peer_alloc(b200_ctx* ctx, size_t bytes, void** out_ptr);
B200_API int b200_peer_free(b200_ctx* ctx, void* ptr);
B200_API int b200_peer_export(b200_ctx* ctx, void* ptr, unsigned char* out_handle /* 64 bytes */);
B200_API int b200_peer_open(b200_ctx* ctx, const unsigned char* handle /* 64 bytes */, void** out_ptr);
B200_API int b200_peer_close(b200_ctx* ctx, void* ptr);

/* ---- Grouper + hash aggregates ------------------------------------------------------------
 * b200_agg_kind: the registered hash aggregate functions (kernels/hash_aggregate.cc:1039-1062)
 * plus hash_mean, which the reference lacks and is defined as MeanImpl::Finalize
 * (kernels/aggregate_basic_internal.h:99-109) over GroupedSumImpl's sums/counts. */
typedef enum b200_agg_kind {
  B200_HASH_COUNT = 0,   /* GroupedCountImpl   hash_aggregate.cc:466-517 */
  B200_HASH_SUM = 1,     /* GroupedSumImpl     hash_aggregate.cc:522-622 */
  B200_HASH_MIN_MAX = 2, /* GroupedMinMaxImpl  hash_aggregate.cc:627-768 */
  B200_HASH_MEAN = 3
} b200_agg_kind;

typedef struct b200_agg_spec {
  int32_t kind;     /* b200_agg_kind */
  int32_t arg_type; /* b200_type of the argument column */
  /* hash_count: CountOptions::count_mode (0 COUNT_NON_NULL, 1 COUNT_NULL; api_aggregate.h:47-58)
   * hash_min_max: MinMaxOptions::null_handling (0 SKIP, 1 EMIT_NULL; api_aggregate.h:65-76) */
  int32_t option;
  int32_t reserved;
} b200_agg_spec;

typedef struct b200_groupby b200_groupby;

/* == Grouper::Make + GetKernels/InitKernels (hash_aggregate.cc:314-356, 809-859).
 * Keys: 1..16 columns, fixed-width (B200_BOOL..B200_DOUBLE) or variable-length (B200_BINARY: binary / utf8 with
 * int32 offsets).  Up to 4 fixed-width columns of <= 64 packed bits are grouped on a packed register key; anything
 * else goes through a key -> id table first (csrc/groupby_wide.cuh).  Other key types: B200_NOT_IMPLEMENTED, as
 * GrouperImpl::Make does ("Keys of type ...", hash_aggregate.cc:352).
 * expected_groups: sizing hint for the open-addressing table (0 = unknown, the table grows). */
B200_API int b200_groupby_create(b200_ctx* ctx, int n_keys, const int32_t* key_types, int n_aggs,
                                 const b200_agg_spec* aggs, int64_t expected_groups,
                                 b200_groupby** out);
B200_API int b200_groupby_destroy(b200_groupby* gb);
/* Bytes the value buffer of variable-length key column `key_index` needs at b200_groupby_finalize (== the sum of
 * the distinct keys' lengths, what VarLengthKeyEncoder::Decode adds up, hash_aggregate.cc:282-285); 0 for a
 * fixed-width key.  Call it after the last consume. */
B200_API int b200_groupby_key_bytes(b200_groupby* gb, int key_index, int64_t* out_bytes);
/* Execution knobs of a handle (no reference counterpart; results never depend on them).  Names:
 *   "hp_min_rows" / "hp_min_groups"   batch size and expected cardinality from which a consume call takes the
 *                                     partition-then-aggregate path (defaults 2^22 rows, 2^21 groups)
 *   "hp_keys_per_part"                distinct keys that path aims at per final partition (default 1536 of the 4096
 *                                     slots of a partition's shared-memory table; clamped to [64, 2048])
 * B200_KEY_ERROR for an unknown name. */
B200_API int b200_groupby_set_option(b200_groupby* gb, const char* name, int64_t value);
/* Which way the handle's consume calls went (no reference counterpart).  Names:
 *   "partitioned_calls"            consume calls served by the partition-then-aggregate path
 *   "partitioned_alloc_failures"   calls that qualified but whose partition buffers (2 x 24 B per row) could not be
 *                                  allocated: served by the global-table kernels instead
 *   "partitioned_span_refusals"    calls with explicit row numbers spanning 2^31 or more: likewise
 * B200_KEY_ERROR for an unknown name. */
B200_API int b200_groupby_stat(b200_groupby* gb, const char* name, int64_t* out);
/* == one iteration of GroupBy's batch loop (hash_aggregate.cc:928-943): Grouper::Consume
 * (:358-408) fused with every aggregate's Consume.  args[i] feeds aggs[i].  row_base is the
 * global index of the batch's first row: group order is first-occurrence order of
 * (row_base + i), which makes range-sharded partial tables mergeable in reference order. */
B200_API int b200_groupby_consume(b200_groupby* gb, const b200_col* keys, const b200_col* args,
                                  uint64_t row_base);
/* As b200_groupby_consume with an explicit global row number per row (row_ids: device uint64[length])
 * instead of row_base + i: the form the hash-repartitioned multi-GPU path uses, where a rank
 * aggregates rows that arrived from every shard (SURVEY §8e, high cardinality). */
B200_API int b200_groupby_consume_rows(b200_groupby* gb, const b200_col* keys, const b200_col* args,
                                       const uint64_t* row_ids);
/* As b200_groupby_consume over the rows whose bit is set in `selection` only: == consuming
 * Filter(column, selection, DROP) of every key and argument column for a selection without nulls, without
 * materialising the filtered columns (compare -> filter -> aggregate in fewer passes; the
 * Q1 shape).  selection: DEVICE bitmap at bit offset 0, ceil(length / 32) whole 32-bit words, e.g. the output
 * of b200_compare_scalar.  Group order is the first-occurrence order of the selected rows, as after a Filter. */
B200_API int b200_groupby_consume_selected(b200_groupby* gb, const b200_col* keys, const b200_col* args,
                                           uint64_t row_base, const void* selection);
/* Compare -> filter -> derive -> aggregate in one call (the Q1 shape of config C5).  Upstream this
 * is four operations over whole columns: a comparison (compute/kernels/scalar_compare.cc) producing the mask, Filter
 * of every used column (vector_selection.cc:1776-1866), "multiply" / "subtract" for the derived measures
 * (scalar_arithmetic.cc:69-190) and GroupBy (hash_aggregate.cc:890-966).  Here:
 *   args[j]    the argument of aggregate j: a column (kind 0), or `a op1 (scalar op2 b)` over float64 columns
 *              (kind 1; e.g. extendedprice * (1 - discount) = {MULTIPLY, SUBTRACT, a = price, scalar = 1, b = discount}),
 *              each operation a single IEEE operation as in b200_arithmetic, validity = a's AND b's
 *   predicate  nullable; rows with `column OP *scalar` false do not exist (== Filter(DROP) of every column by the
 *              comparison's result).  The predicate column must hold no nulls.
 * The result is that of the four operations run one after the other (group order included).  For <= 8 expected
 * groups, byte-wide keys and float64 arguments without nulls every input byte is read exactly once
 * (gb_fused8_kernel); other shapes are evaluated through b200_compare_scalar / b200_arithmetic /
 * b200_groupby_consume_selected internally. */
typedef struct b200_fused_arg {
  int32_t kind;     /* 0 = column a; 1 = a op1 (scalar op2 b) */
  int32_t op1, op2; /* b200_arith_op */
  int32_t reserved;
  b200_col a, b;
  double scalar;
} b200_fused_arg;
typedef struct b200_predicate {
  b200_col column;
  int32_t op;          /* 0 equal, 1 not_equal, 2 greater, 3 greater_equal, 4 less, 5 less_equal */
  int32_t reserved;
  const void* scalar;  /* HOST pointer to one value of the column's type */
} b200_predicate;
B200_API int b200_groupby_consume_fused(b200_groupby* gb, const b200_col* keys, const b200_fused_arg* args,
                                        uint64_t row_base, const b200_predicate* predicate /* nullable */);
/* global first-occurrence row of every group, in output (group id) order: device uint64[num_groups].
 * Lets ranks holding disjoint key sets restore the reference's global group order
 * (hash_aggregate.cc:391-403) by ranking their groups on it. */
B200_API int b200_groupby_first_rows(b200_groupby* gb, uint64_t* out_first_rows);
/* == Grouper::Consume proper (hash_aggregate.cc:358-408): writes the uint32 group id of every
 * row, ids handed out in first-occurrence order across calls. */
B200_API int b200_grouper_consume(b200_groupby* gb, const b200_col* keys, uint32_t* out_ids);
/* == Grouper::MakeGroupings (hash_aggregate.cc:968-1001; api_aggregate.h:333-359): for every group id in
 * [0, num_groups) the ascending list of rows holding it, as the two buffers of a list<int32> array:
 * out_offsets int32[num_groups + 1], out_indices int32[length] (both DEVICE).  Used upstream by dataset partitioning
 * (dataset/partition.cc:115-122).  Grouper::ApplyGroupings (hash_aggregate.cc:1003-1018) is then
 * b200_take(array, out_indices) under the same offsets.  An id >= num_groups fails with B200_INDEX_ERROR.
 * Synchronises. */
B200_API int b200_make_groupings_scratch_bytes(int64_t length, size_t* out_bytes);
B200_API int b200_make_groupings(b200_ctx* ctx, const uint32_t* ids, int64_t length, uint32_t num_groups,
                                 int32_t* out_offsets, int32_t* out_indices, void* scratch, size_t scratch_bytes);
/* == DictEncodeAction (compute/kernels/vector_hash.cc:165-226), the indices half of "dictionary_encode": the
 * uint32 group ids b200_grouper_consume wrote for one key column -> int32 dictionary indices.
 * null_group < 0: the column has no null, or DictionaryEncodeOptions::ENCODE (the null keeps its dictionary entry):
 *   indices = ids, no validity written (out_validity may be NULL).
 * null_group >= 0 (DictionaryEncodeOptions::MASK, the default): the id of the null key's group, which has NO
 *   dictionary entry: its rows get a null index (value 0), ids above it shift down by one, and out_validity
 *   (DEVICE, ceil(length / 32) whole 32-bit words) receives the index validity.  The dictionary is the handle's
 *   key column (b200_groupby_finalize) with entry null_group removed.  Asynchronous on the context's stream. */
B200_API int b200_dictionary_indices(b200_ctx* ctx, const uint32_t* ids, int64_t length, int64_t null_group,
                                     int32_t* out_indices, void* out_validity);
/* == HashAggregateKernel::consume with the reference signature (arg, uint32 ids, num_groups)
 * (hash_aggregate.cc:783-789): accumulates aggregate `agg_index` of `gb` by dense group id. */
B200_API int b200_hash_agg_consume(b200_groupby* gb, int agg_index, const b200_col* arg,
                                   const uint32_t* ids, uint32_t num_groups);
/* == HashAggregateKernel::merge (kernel.h:688-690; NotImplemented upstream, hash_aggregate.cc:795-798)
 * for two handles driven through b200_hash_agg_consume over the same dense id space: folds every
 * group state of `src` into `dst` slot by slot (same aggregate schema). */
B200_API int b200_hash_agg_merge(b200_groupby* dst, b200_groupby* src);
/* == Grouper::num_groups (hash_aggregate.cc:410); synchronises. */
B200_API int b200_groupby_num_groups(b200_groupby* gb, uint32_t* out /* host */);
/* == HashAggregateKernel::merge, which the reference leaves NotImplemented
 * (hash_aggregate.cc:795-798): folds every group of `src` into `dst` (same key/agg schema). */
B200_API int b200_groupby_merge(b200_groupby* dst, b200_groupby* src);
/* Export / import of the occupied slots as a flat device buffer, the unit the multi-GPU
 * all-gather exchanges.  state_bytes = num_groups * slot_bytes. */
B200_API int b200_groupby_state_layout(b200_groupby* gb, size_t* slot_bytes);
B200_API int b200_groupby_export(b200_groupby* gb, void* out_state, uint32_t* out_groups /* host */);
B200_API int b200_groupby_import(b200_groupby* gb, const void* state, uint32_t n_slots);
/* == finalize of every kernel + Grouper::GetUniques (hash_aggregate.cc:946-960, 412-429).
 * Call num_groups first, allocate, then finalize.  Layout of the outputs (all of length G):
 *   out_keys[k]           data = G elements of key type k; validity = G bits or NULL
 *                         (library reports null_count; validity may be NULL only if the key
 *                         column never carried a validity bitmap)
 *   out_aggs[j], by kind  COUNT:   data int64[G], validity ignored (never null)
 *                         SUM:     data {int64|uint64|double}[G], validity G bits (required)
 *                         MEAN:    data double[G], validity G bits (required)
 *                         MIN_MAX: TWO consecutive entries: mins then maxes, of the argument
 *                                  type, sharing one validity (the first's; the second's is
 *                                  ignored), like struct<min,max> (hash_aggregate.cc:741-757)
 * so out_aggs has n_aggs + (number of MIN_MAX aggs) entries. */
B200_API int b200_groupby_finalize(b200_groupby* gb, b200_out_col* out_aggs, b200_out_col* out_keys);


/* ---- hash repartitioning (multi-GPU grouped aggregation at high cardinality) -----
 * No reference counterpart: upstream has no distributed execution; this realises north_star's
 * "hash-repartitioning alltoall".  dest(row) = hash(packed key of gb's key schema) % n_parts;
 * out_pos[row] = position of the row in a layout where the rows of destination p are contiguous
 * and destination p starts at sum(out_counts[0..p)).  n_parts <= 16; length < 2^32.
 *   plan     -> out_pos (device uint32[length]), out_counts (host int64[n_parts]); synchronises
 *   scatter  -> out_data[out_pos[i]] = column[i]; out_valid (nullable) gets one 0/1 byte per row
 *   rows     -> out_rows[out_pos[i]] = row_base + i
 *   pack     -> validity bytes (as received from the exchange) to an Arrow bitmap + null count */
B200_API int b200_partition_scratch_bytes(int64_t length, size_t* out_bytes);
B200_API int b200_groupby_partition_plan(b200_groupby* gb, const b200_col* keys, int n_parts,
                                         uint32_t* out_pos, int64_t* out_counts /* host */,
                                         void* scratch, size_t scratch_bytes);
B200_API int b200_partition_scatter(b200_ctx* ctx, const b200_col* col, const uint32_t* pos,
                                    void* out_data, uint8_t* out_valid /* nullable */);
B200_API int b200_partition_rows(b200_ctx* ctx, int64_t length, uint64_t row_base,
                                 const uint32_t* pos, uint64_t* out_rows);
B200_API int b200_pack_validity(b200_ctx* ctx, const uint8_t* valid_bytes, int64_t length,
                                void* out_bitmap, int64_t* out_null_count /* host, nullable */);

#ifdef __cplusplus
}
#endif
#endif /* ARROW1_B200_H */
