/*
 * hsgpu.h — C ABI of libhsgpu, the B200 (sm_100a) covering-index engine that
 * stands behind the Hyperspace API for its one data-parallel hot path:
 * index build (Murmur3 bucket -> scatter -> per-bucket sort), bucketed
 * sort-merge join, filtered index scan.
 *
 * The reference (pirz/hyperspace-1, Scala on Spark 2.4.2) has no FFI seam at the
 * compute level; each entry point below names the reference call site whose
 * Spark-side work it replaces (paths relative to the reference root,
 * HS/ = src/main/scala/com/microsoft/hyperspace/).
 *
 * Conventions
 *  - every function returns an hs_status (0 = HS_OK); the text of the last error
 *    of a context is read with hs_last_error().
 *  - column buffers follow the Arrow columnar layout (LSB validity bitmap,
 *    int32 offsets for UTF8/BINARY, values little-endian, bit offset 0).
 *  - input buffers are BORROWED for the duration of the call and may live in
 *    host memory (pageable or pinned) or in device memory of the context's GPU;
 *    the library looks at the pointer to find out which.
 *  - outputs are hs_table objects owned by the library and released with
 *    hs_table_free(); their buffers are pinned host memory or device memory,
 *    selected by hs_ctx_set_output_location().
 *  - calls are synchronous: the context's stream is drained before returning. The context
 *    launches on its OWN stream: device inputs must be complete before the call (synchronise
 *    the stream that produced them first).
 *  - a context is bound to one device and is not thread-safe; different
 *    contexts may be used concurrently from different threads.
 *  - there is no CPU fallback: unsupported inputs return HS_EUNSUPPORTED.
 */
#ifndef HSGPU_H_
#define HSGPU_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum hs_status {
  HS_OK = 0,
  HS_EINVAL = 1,
  HS_ENOMEM = 2,
  HS_ECUDA = 3,
  HS_ENCCL = 4,
  HS_EUNSUPPORTED = 5
} hs_status;

/* Logical column types (Spark SQL type -> code). DATE is carried as HS_INT32,
 * TIMESTAMP as HS_INT64 (that is how Spark hashes and orders them). */
typedef enum hs_type {
  HS_BOOL = 1,    /* bit-packed, LSB first (Arrow boolean) */
  HS_INT8 = 2,
  HS_INT16 = 3,
  HS_INT32 = 4,
  HS_INT64 = 5,
  HS_FLOAT32 = 6,
  HS_FLOAT64 = 7,
  HS_UTF8 = 8,    /* int32 offsets[length+1] + bytes */
  HS_BINARY = 9,  /* same layout as HS_UTF8 */
  HS_LARGE_UTF8 = 10,   /* Arrow large_string: `offsets` points at int64 entries (cast the pointer); for columns
                         * holding more than 2 GiB of bytes. Hashes and orders exactly like HS_UTF8. */
  HS_LARGE_BINARY = 11  /* same layout as HS_LARGE_UTF8 */
} hs_type;

typedef enum hs_mem { HS_MEM_HOST = 0, HS_MEM_DEVICE = 1 } hs_mem;

/* One Arrow-layout column. `validity` may be NULL when null_count == 0. */
typedef struct hs_column {
  int32_t type;            /* hs_type */
  int32_t reserved;
  int64_t length;          /* rows */
  int64_t null_count;      /* 0 when validity == NULL */
  const uint8_t* validity; /* LSB bitmap, bit i set = row i valid; or NULL */
  const int32_t* offsets;  /* HS_UTF8/HS_BINARY: length+1 int32 entries; HS_LARGE_*: length+1 int64 entries
                            * behind the same pointer */
  const void* data;        /* values (fixed width) or bytes (UTF8/BINARY) */
} hs_column;

/* Library-owned result: columns plus, for indexes and joins, the bucket
 * boundaries (row r of bucket b lives at bucket_offsets[b] <= r <
 * bucket_offsets[b+1]). bucket_offsets is always host memory. */
typedef struct hs_table {
  int32_t ncols;
  int32_t location;              /* hs_mem of the column buffers */
  int64_t nrows;
  hs_column* cols;
  int32_t num_buckets;           /* 0 when the table is not bucketed */
  int32_t reserved;
  const int64_t* bucket_offsets; /* num_buckets+1 entries or NULL */
  void* impl;                    /* private */
} hs_table;

/* A borrowed view of one sorted run per bucket: the in-memory form of the files
 * of one index version (HS/index/IndexLogEntry.scala:507-511 bucketSpec:
 * bucket columns == sort columns == the first n_indexed columns). */
typedef struct hs_index_view {
  const hs_column* cols;
  int32_t ncols;
  int32_t n_indexed;
  int32_t num_buckets;
  int32_t reserved;
  const int64_t* bucket_offsets; /* host memory, num_buckets+1 entries */
} hs_index_view;

/* ---- filter predicates (postfix program, SQL three-valued logic) ---------- */
typedef enum hs_pred_op {
  HS_PRED_EQ = 1, HS_PRED_NE = 2, HS_PRED_LT = 3, HS_PRED_LE = 4,
  HS_PRED_GT = 5, HS_PRED_GE = 6,          /* column <op> literal          */
  HS_PRED_IS_NULL = 7, HS_PRED_IS_NOT_NULL = 8,
  HS_PRED_IN = 9,  /* integer column IN (list): lit_bytes points at lit_len int64 values (host memory, any order).
                    * NOT IN = IN followed by NOT. This is the lineage filter of hybrid scan and of an incremental
                    * refresh after source files were deleted: NOT (_data_file_id IN (ids of the deleted files)),
                    * HS/index/rules/RuleUtils.scala:402-425, HS/actions/RefreshIncrementalAction.scala:73-95 */
  HS_PRED_AND = 16, HS_PRED_OR = 17, HS_PRED_NOT = 18 /* pop operands      */
} hs_pred_op;

typedef struct hs_pred_term {
  int32_t op;            /* hs_pred_op */
  int32_t col;           /* column index in the view (leaf ops) */
  int64_t lit_i64;       /* literal for BOOL/INT8..INT64 columns */
  double lit_f64;        /* literal for FLOAT32/FLOAT64 columns */
  const char* lit_bytes; /* literal for UTF8/BINARY columns; the int64 list of HS_PRED_IN (host memory) */
  int32_t lit_len;       /* bytes of the string literal; number of values of HS_PRED_IN */
  int32_t reserved;
} hs_pred_term;

typedef struct hs_predicate {
  int32_t nterms;
  int32_t reserved;
  const hs_pred_term* terms; /* postfix order; at most 32 terms, depth 16 */
} hs_predicate;

/* JoinIndexRule matches a Join of any type with an equi-condition (HS/index/rules/JoinIndexRule.scala:57-60);
 * these are the types Spark's SortMergeJoinExec runs over the two bucketed index scans. Rows whose key holds a
 * NULL match nothing: they appear (padded with NULLs) in the outer joins on their own side and in LEFT_ANTI.
 * SEMI and ANTI return left columns only (n_right_out must be 0). */
typedef enum hs_join_type {
  HS_JOIN_INNER = 0,
  HS_JOIN_LEFT_OUTER = 1,
  HS_JOIN_RIGHT_OUTER = 2,
  HS_JOIN_FULL_OUTER = 3,
  HS_JOIN_LEFT_SEMI = 4,
  HS_JOIN_LEFT_ANTI = 5
} hs_join_type;

/* Per-stage device timings of the last call (CUDA events on the context's
 * stream), filled only while timing is enabled. */
typedef struct hs_stage_timing {
  char name[32];
  float ms;          /* elapsed device time of the stage */
  int32_t launches;  /* kernels launched by the stage */
} hs_stage_timing;

typedef struct hs_ctx hs_ctx;

/* ---- context ------------------------------------------------------------- */
/* One context per (thread, device); mirrors the reference's thread-local
 * HyperspaceContext (HS/Hyperspace.scala:168-204). */
int32_t hs_ctx_create(int32_t device, hs_ctx** out);
int32_t hs_ctx_destroy(hs_ctx* ctx);
const char* hs_last_error(hs_ctx* ctx);
/* HS_MEM_HOST (default, pinned) or HS_MEM_DEVICE for the buffers of later results. */
int32_t hs_ctx_set_output_location(hs_ctx* ctx, int32_t location);
/* Deferred host results (off by default). With it on, a call that returns an hs_table in host memory comes
 * back as soon as the table's device -> host copies are queued on a second stream of the context; the table's
 * buffers may be read only after hs_table_wait() (hs_table_free() waits by itself). One caller thread can then
 * start the next build — whose input upload runs beside the previous result's download — and collect the
 * previous table afterwards. This is the one exception to "calls are synchronous"; it mirrors how the
 * reference's driver thread hands a finished stage to Spark's writer while the next stage is already running
 * (CreateActionBase.write, HS/actions/CreateActionBase.scala:122-140, is one call per index but many per job). */
int32_t hs_ctx_set_deferred_results(hs_ctx* ctx, int32_t on);
int32_t hs_table_wait(hs_ctx* ctx, hs_table* table);
/* Memory kept between calls. A context allocates from a stream-ordered pool of its own (the device's default
 * pool is not touched) and keeps the large device blocks and the pinned result buffers it has freed for the next
 * call, up to half of the device memory and 16 GiB of pinned memory (environment: HSGPU_CACHE_MB,
 * HSGPU_PINNED_CACHE_MB; the blocks freed longest ago are released beyond that). hs_ctx_trim() gives everything
 * cached back to the driver, for hosts that share the GPU with other libraries. */
int32_t hs_ctx_trim(hs_ctx* ctx);
/* The cudaStream_t the context launches on (for external event timing). */
int32_t hs_ctx_get_stream(hs_ctx* ctx, void** stream);
int32_t hs_ctx_enable_timing(hs_ctx* ctx, int32_t on);
int32_t hs_ctx_last_timings(hs_ctx* ctx, hs_stage_timing* out, int32_t cap, int32_t* n);
/* Kernels launched by this context since creation. */
int32_t hs_ctx_launch_count(hs_ctx* ctx, int64_t* out);
int32_t hs_table_free(hs_ctx* ctx, hs_table* t);
const char* hs_version(void);

/* ---- K1: bucket id of every row ------------------------------------------
 * Replaces Spark's HashPartitioning(cols, n).partitionIdExpression =
 * Pmod(Murmur3Hash(cols, 42), n), reached from
 * df.repartition(numBuckets, indexedCols) at HS/actions/CreateActionBase.scala:129-130
 * (again at HS/actions/OptimizeAction.scala:90-91, HS/index/rules/RuleUtils.scala:574)
 * and, inside the writer, from BucketSpec at HS/index/DataFrameWriterExtensions.scala:61-63.
 * out_bucket_ids: keys[0].length int32 values, host or device memory. */
int32_t hs_hash_bucket(hs_ctx* ctx, const hs_column* keys, int32_t nkeys,
                       int32_t num_buckets, int32_t* out_bucket_ids);
/* Raw Murmur3Hash(cols, 42) per row (the parity probe for Appendix A.6 vectors). */
int32_t hs_hash_rows(hs_ctx* ctx, const hs_column* keys, int32_t nkeys,
                     int32_t* out_hashes);

/* Bucket ids and per-bucket row counts for callers that cannot allocate GPU memory.
 * No counterpart in the reference (it never touches a GPU); it exists for hosts where cudaMalloc is
 * not an option: an executor whose device memory is wholly reserved by another pool (an arena /
 * RMM-style allocator sized to the GPU), or a thread that may not trigger the device-wide
 * synchronisation cudaMalloc/cudaFree imply. The histogram counters live in a fixed 32 KiB array
 * inside the library image, so num_buckets is limited to HS_NOALLOC_MAX_BUCKETS = 8192 (larger
 * values return HS_EINVAL), and its counters are 32 bits wide, so a call takes fewer than 2^32-16
 * rows (more returns HS_EUNSUPPORTED). Key buffers and out_bucket_ids (int32 per row) must be device memory;
 * out_counts (num_buckets uint32) may be host or device memory. All calls on a GPU share the
 * array: do not let two of them overlap. The first call on a device resolves the array's device
 * address; make that call outside of stream capture. Same bucket ids as hs_hash_bucket. */
#define HS_NOALLOC_MAX_BUCKETS 8192
int32_t hs_bucket_histogram_noalloc(hs_ctx* ctx, const hs_column* keys, int32_t nkeys,
                                    int32_t num_buckets, int32_t* out_bucket_ids,
                                    uint32_t* out_counts);

/* ---- K1+K2+K3: the index build --------------------------------------------
 * Replaces the body of CreateActionBase.write (HS/actions/CreateActionBase.scala:122-140):
 * repartition(numBuckets, indexedCols) followed by saveWithBuckets with
 * BucketSpec(n, cols, cols) (HS/index/DataFrameWriterExtensions.scala:49-67), i.e.
 * hash-partition, then sort every bucket on the indexed columns ASC NULLS FIRST.
 * The first n_indexed columns are the indexed (bucket == sort) columns, the rest
 * are included columns. Result: all columns in bucket-contiguous, per-bucket
 * sorted order + bucket_offsets. Parquet encoding is the caller's job (a5). */
int32_t hs_build_index(hs_ctx* ctx, const hs_column* cols, int32_t ncols,
                       int32_t n_indexed, int32_t num_buckets, hs_table** out);

/* ---- compaction / multi-file buckets --------------------------------------
 * Merges nruns index runs with equal schema and bucket count into one run per
 * bucket. Replaces the repartition + saveWithBuckets of OptimizeAction.op
 * (HS/actions/OptimizeAction.scala:85-99) and the per-side SortExec Spark plans
 * when a bucket has more than one file after an incremental refresh
 * (HS/actions/RefreshIncrementalAction.scala:129-144; asserted in
 * src/test/scala/com/microsoft/hyperspace/index/E2EHyperspaceRulesTest.scala:474-478).
 * Every run must be what hs_build_index returns for its rows: rows of bucket b at
 * [bucket_offsets[b], bucket_offsets[b+1]), sorted on the indexed columns. Fixed-width tables
 * with NULL-free keys are merged along the merge path (no sort); other shapes are
 * concatenated and sorted. */
int32_t hs_merge_runs(hs_ctx* ctx, const hs_index_view* runs, int32_t nruns,
                      hs_table** out);

/* ---- K5+K6: bucketed sort-merge join, no exchange --------------------------
 * Executes the plan JoinIndexRule.apply produces (HS/index/rules/JoinIndexRule.scala:57-90)
 * over two index relations created by RuleUtils.transformPlanToUseIndexOnlyScan with
 * useBucketSpec = true (HS/index/rules/RuleUtils.scala:264-292): bucket i of the left
 * index joins bucket i of the right one on the indexed columns; NULL keys never match.
 * Key types must agree (else HS_EUNSUPPORTED). If the bucket counts differ (the rule tolerates it,
 * JoinIndexRule.scala:515-516; Spark re-shuffles one side), the side with fewer rows is re-bucketed
 * on the fly to the other side's count and the result carries that count.
 * Output columns: left_out[0..n_left_out) of the left view, then right_out[..] of the
 * right view; row order inside a bucket is unspecified (compare as multisets). In the outer
 * joins the columns of the side without a match are NULL for that row (validity bitmaps are
 * added to the result as needed). join_type: hs_join_type. */
int32_t hs_join_bucketed(hs_ctx* ctx, const hs_index_view* left,
                         const hs_index_view* right, int32_t join_type,
                         const int32_t* left_out, int32_t n_left_out,
                         const int32_t* right_out, int32_t n_right_out,
                         hs_table** out);

/* ---- K4: Parquet page bodies (PLAIN) --------------------------------------
 * What Spark's DynamicPartitionDataWriter + ParquetFileFormat do to every value behind saveWithBuckets
 * (HS/index/DataFrameWriterExtensions.scala:58-65). One call encodes all data pages (Parquet v1, PLAIN,
 * uncompressed) of ONE column: page p covers rows [page_rows[p], page_rows[p+1]); page_rows (host memory) has
 * npages+1 entries, starts at 0 and ends at the column length. Result: column 0 = BINARY with one row per page,
 * the page body = [definition levels, only if `optional`] ++ [PLAIN values of the non-NULL rows]; column 1 =
 * INT64 number of NULLs per page. Definition levels: 4-byte LE byte count + RLE/bit-packed hybrid of bit
 * width 1 (one RLE run for a page without NULLs, bit-packed groups from the validity bitmap otherwise).
 * Values: fixed width little-endian; BOOL bit-packed LSB first; UTF8/BINARY 4-byte LE length + bytes.
 * The Thrift page headers and the file footer are framing the host writes. */
int32_t hs_parquet_plain_pages(hs_ctx* ctx, const hs_column* col, int32_t optional,
                               const int64_t* page_rows, int32_t npages, hs_table** out);

/* The same with parquet-mr's first choice of encoding, the per-chunk dictionary (what Spark 2.4 writes for the
 * index files: ParquetOutputFormat, dictionary enabled): page_chunk[p] (host, non-decreasing, < nchunks) names
 * the column chunk — here: the bucket file — of page p. A chunk of a 4- or 8-byte column whose distinct non-NULL
 * values number at most max_dict (<= 32768) gets a dictionary: its data pages are [definition levels] ++ [bit
 * width byte] ++ [one bit-packed run of the indices] (encoding PLAIN_DICTIONARY), and its dictionary page body
 * (PLAIN values in index order) is row npages + chunk of column 0. All other chunks (too many distinct values,
 * strings, booleans, max_dict = 0) are PLAIN exactly as above and their dictionary row is empty. Column 1: NULLs
 * per page for rows < npages, dictionary entries (0 = PLAIN chunk) for rows npages .. npages + nchunks - 1.
 * codec: 0 UNCOMPRESSED, or 1 SNAPPY (Spark's default, spark.sql.parquet.compression.codec): every body of
 * column 0 is then the raw snappy stream of the page (length preamble, literals and copies; any snappy
 * decompressor reads it), and a third column (INT64) gives the uncompressed size the page header needs. */
int32_t hs_parquet_encode_pages(hs_ctx* ctx, const hs_column* col, int32_t optional,
                                const int64_t* page_rows, int32_t npages, const int32_t* page_chunk,
                                int32_t nchunks, int32_t max_dict, int32_t codec, hs_table** out);

/* ---- K5: Parquet pages decoded on the device ---------------------------------
 * The read side of the same files: what Spark's ParquetFileFormat reader does under the bucketed scan the rules
 * plan (HS/index/rules/RuleUtils.scala:280-286). One call decodes ONE column over all the pages handed in (any
 * number of files / column chunks, in row order) into one column: codec UNCOMPRESSED (0) or SNAPPY (1); data
 * pages v1 of flat columns (definition levels: RLE/bit-packed hybrid, bit width 1, only when `optional`); values
 * PLAIN or dictionary-encoded (PLAIN_DICTIONARY 2 / RLE_DICTIONARY 8, with the chunk's PLAIN dictionary page
 * among the pages). `bytes` (host or device) holds the page bodies as they sit in the files, framing excluded;
 * the host parses footers and page headers (hyperspace-1_b200/parquet_read.py). `type`: the hs_type wanted
 * (HS_INT8 / HS_INT16 are narrowed from Parquet's INT32). Other codecs, encodings and page kinds return
 * HS_EUNSUPPORTED. */
typedef struct hs_parquet_page {
  int64_t offset;             /* of the page body inside `bytes` */
  int64_t size;               /* bytes of the body as stored (compressed when codec != 0) */
  int64_t uncompressed_size;
  int64_t num_values;         /* rows of a data page (NULLs included); entries of a dictionary page */
  int32_t kind;               /* 0 data page (v1), 2 dictionary page */
  int32_t encoding;           /* of the values: 0 PLAIN, 2 PLAIN_DICTIONARY, 8 RLE_DICTIONARY */
  int32_t chunk;              /* column chunk the page belongs to: the scope of its dictionary */
  int32_t reserved;
} hs_parquet_page;
int32_t hs_parquet_decode_pages(hs_ctx* ctx, const uint8_t* bytes, int64_t nbytes, const hs_parquet_page* pages,
                                int32_t npages, int32_t type, int32_t optional, int32_t codec, hs_table** out);

/* ---- K7: filtered index scan -----------------------------------------------
 * Executes the plan FilterIndexRule.apply produces (HS/index/rules/FilterIndexRule.scala:43-85,
 * useBucketSpec = false at :59-65): rows of the index for which the predicate is TRUE,
 * projected to out_cols. When the predicate is `first indexed column = literal` on a
 * single-column index the scan prunes to the literal's bucket and binary-searches the
 * sorted run; the result multiset is the same. */
int32_t hs_filter_scan(hs_ctx* ctx, const hs_index_view* index,
                       const hs_predicate* pred, const int32_t* out_cols,
                       int32_t n_out, hs_table** out);

/* ---- multi-GPU build (one process per GPU) ----------------------------------
 * Replaces Spark's ShuffleExchangeExec under the same repartition call
 * (HS/actions/CreateActionBase.scala:129-130): rows travel once to the GPU that owns
 * their bucket, owner(b) = b * nranks / num_buckets: the scatter kernel stores them straight into
 * the owner's peer-mapped memory over NVLink (NCCL sets up, synchronises and is the fallback when peers
 * cannot be mapped); string columns and keys wider than 64 bits change hands as whole rows in one
 * grouped ncclSend/ncclRecv round before a local build. Every later step is bucket-local. hs_comm_unique_id fills a 128-byte ncclUniqueId
 * on rank 0 which the host passes to every rank (any transport). */
int32_t hs_comm_unique_id(hs_ctx* ctx, void* out_id128);
int32_t hs_comm_init(hs_ctx* ctx, int32_t rank, int32_t nranks, const void* id128);
int32_t hs_comm_destroy(hs_ctx* ctx);
/* As hs_build_index, on this rank's share of the input rows. The result holds the
 * buckets this rank owns (other buckets are empty ranges in bucket_offsets). */
int32_t hs_build_index_dist(hs_ctx* ctx, const hs_column* cols, int32_t ncols,
                            int32_t n_indexed, int32_t num_buckets, hs_table** out);
/* Bytes this rank sent to other ranks and the device time of the exchange in the last
 * hs_build_index_dist call (for the all-to-all bus-bandwidth figure). */
int32_t hs_comm_last_exchange(hs_ctx* ctx, int64_t* bytes_sent, float* ms);

#ifdef __cplusplus
}
#endif
#endif /* HSGPU_H_ */
