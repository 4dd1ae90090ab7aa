/* hsgpu_jni.c — the JNI shim between the reference's Scala host code and libhsgpu (include/hsgpu.h).
 *
 * One function per export. Columns cross the boundary as a direct ByteBuffer holding an array of hs_column
 * structs (48 bytes each, little-endian) whose pointers are the addresses of Arrow off-heap buffers
 * (ArrowBuf.memoryAddress()); Spark 2.4.2 bundles Arrow 0.10, which predates the C Data Interface, so raw
 * addresses are what the JVM side has. A non-zero hs_status becomes a
 * com.microsoft.hyperspace.HyperspaceException carrying hs_last_error() — the exception
 * Action.run (HS/actions/Action.scala:96-103) logs and rethrows and the rules swallow
 * (HS/index/rules/JoinIndexRule.scala:85-89, FilterIndexRule.scala:79-83).
 *
 * Build (needs a JDK; none in the development image, so the Makefile beside this file skips it there):
 *   gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -I../../include hsgpu_jni.c \
 *       -L../../hyperspace-1_b200 -lhsgpu -o libhsgpu_jni.so
 * tests/test_bindings_cpu.py compiles this file with -fsyntax-only against a minimal jni.h stand-in, so that
 * every call into the C ABI is type-checked against include/hsgpu.h on every run of the CPU tests.
 */
#include <jni.h>
#include <stdint.h>
#include <string.h>

#include "hsgpu.h"

#define HS_JNI(ret, name) JNIEXPORT ret JNICALL Java_com_microsoft_hyperspace_gpu_HsGpu_##name

static hs_ctx* as_ctx(jlong h) { return (hs_ctx*)(intptr_t)h; }
static hs_table* as_table(jlong h) { return (hs_table*)(intptr_t)h; }

/* hs_status != 0 -> HyperspaceException(message) (HS/HyperspaceException.scala:19) */
static void throw_hs(JNIEnv* env, hs_ctx* ctx, int32_t st) {
  static const char* names[] = {"HS_OK", "HS_EINVAL", "HS_ENOMEM", "HS_ECUDA", "HS_ENCCL", "HS_EUNSUPPORTED"};
  char msg[640];
  const char* detail = ctx ? hs_last_error(ctx) : "no context";
  const char* nm = (st >= 0 && st <= 5) ? names[st] : "HS_?";
  size_t n = strlen(nm);
  memcpy(msg, nm, n);
  msg[n++] = ':';
  msg[n++] = ' ';
  strncpy(msg + n, detail ? detail : "", sizeof msg - n - 1);
  msg[sizeof msg - 1] = 0;
  jclass cls = (*env)->FindClass(env, "com/microsoft/hyperspace/HyperspaceException");
  if (cls) (*env)->ThrowNew(env, cls, msg);
}

/* a direct ByteBuffer (or null) -> its address */
static const void* buf_addr(JNIEnv* env, jobject buf) {
  return buf ? (*env)->GetDirectBufferAddress(env, buf) : NULL;
}

/* ---- context: one per (thread, device), mirroring the thread-local HyperspaceContext
 * (HS/Hyperspace.scala:168-204) */
HS_JNI(jlong, ctxCreate)(JNIEnv* env, jclass cls, jint device) {
  (void)cls;
  hs_ctx* ctx = NULL;
  int32_t st = hs_ctx_create(device, &ctx);
  if (st) { throw_hs(env, NULL, st); return 0; }
  return (jlong)(intptr_t)ctx;
}

HS_JNI(void, ctxDestroy)(JNIEnv* env, jclass cls, jlong ctx) {
  (void)env; (void)cls;
  hs_ctx_destroy(as_ctx(ctx));
}

HS_JNI(void, ctxSetOutputLocation)(JNIEnv* env, jclass cls, jlong ctx, jint location) {
  (void)cls;
  int32_t st = hs_ctx_set_output_location(as_ctx(ctx), location);
  if (st) throw_hs(env, as_ctx(ctx), st);
}

HS_JNI(void, ctxSetDeferredResults)(JNIEnv* env, jclass cls, jlong ctx, jboolean on) {
  (void)cls;
  int32_t st = hs_ctx_set_deferred_results(as_ctx(ctx), on ? 1 : 0);
  if (st) throw_hs(env, as_ctx(ctx), st);
}

/* ---- K1: bucket ids, the parity probe of HashPartitioning.partitionIdExpression
 * (HS/actions/CreateActionBase.scala:129-130). outIds: direct IntBuffer of `rows` entries. */
HS_JNI(void, hashBucket)(JNIEnv* env, jclass cls, jlong ctx, jobject keys, jint nkeys, jint numBuckets, jobject outIds) {
  (void)cls;
  int32_t st = hs_hash_bucket(as_ctx(ctx), (const hs_column*)buf_addr(env, keys), nkeys, numBuckets,
                              (int32_t*)buf_addr(env, outIds));
  if (st) throw_hs(env, as_ctx(ctx), st);
}

/* ---- the build: replaces the body of CreateActionBase.write (HS/actions/CreateActionBase.scala:122-140) */
HS_JNI(jlong, buildIndex)(JNIEnv* env, jclass cls, jlong ctx, jobject cols, jint ncols, jint nIndexed, jint numBuckets) {
  (void)cls;
  hs_table* out = NULL;
  int32_t st = hs_build_index(as_ctx(ctx), (const hs_column*)buf_addr(env, cols), ncols, nIndexed, numBuckets, &out);
  if (st) { throw_hs(env, as_ctx(ctx), st); return 0; }
  return (jlong)(intptr_t)out;
}

/* one process (executor task) per GPU: HS/actions/CreateActionBase.scala:129-130 at N executors */
HS_JNI(jbyteArray, commUniqueId)(JNIEnv* env, jclass cls, jlong ctx) {
  (void)cls;
  jbyte id[128];
  int32_t st = hs_comm_unique_id(as_ctx(ctx), id);
  if (st) { throw_hs(env, as_ctx(ctx), st); return NULL; }
  jbyteArray a = (*env)->NewByteArray(env, 128);
  if (a) (*env)->SetByteArrayRegion(env, a, 0, 128, id);
  return a;
}

HS_JNI(void, commInit)(JNIEnv* env, jclass cls, jlong ctx, jint rank, jint nranks, jbyteArray id128) {
  (void)cls;
  jbyte id[128];
  (*env)->GetByteArrayRegion(env, id128, 0, 128, id);
  int32_t st = hs_comm_init(as_ctx(ctx), rank, nranks, id);
  if (st) throw_hs(env, as_ctx(ctx), st);
}

HS_JNI(void, commDestroy)(JNIEnv* env, jclass cls, jlong ctx) {
  (void)env; (void)cls;
  hs_comm_destroy(as_ctx(ctx));
}

HS_JNI(jlong, buildIndexDist)(JNIEnv* env, jclass cls, jlong ctx, jobject cols, jint ncols, jint nIndexed, jint numBuckets) {
  (void)cls;
  hs_table* out = NULL;
  int32_t st = hs_build_index_dist(as_ctx(ctx), (const hs_column*)buf_addr(env, cols), ncols, nIndexed, numBuckets, &out);
  if (st) { throw_hs(env, as_ctx(ctx), st); return 0; }
  return (jlong)(intptr_t)out;
}

/* ---- index views: {hs_column[] buffer, ncols, nIndexed, numBuckets, long[] bucketOffsets} ----------------- */
typedef struct view_hold {
  hs_index_view v;
  jlongArray arr;
  jlong* off;
} view_hold;

static int view_open(JNIEnv* env, view_hold* h, jobject cols, jint ncols, jint nIndexed, jint numBuckets, jlongArray off) {
  h->arr = off;
  h->off = (*env)->GetLongArrayElements(env, off, NULL);
  if (!h->off) return 0;
  h->v.cols = (const hs_column*)buf_addr(env, cols);
  h->v.ncols = ncols;
  h->v.n_indexed = nIndexed;
  h->v.num_buckets = numBuckets;
  h->v.reserved = 0;
  h->v.bucket_offsets = (const int64_t*)h->off;   /* jlong is int64_t */
  return 1;
}
static void view_close(JNIEnv* env, view_hold* h) {
  if (h->off) (*env)->ReleaseLongArrayElements(env, h->arr, h->off, JNI_ABORT);
}

/* ---- compaction and multi-file buckets: OptimizeAction.op (HS/actions/OptimizeAction.scala:85-99),
 * RefreshIncrementalAction (HS/actions/RefreshIncrementalAction.scala:129-144). Two runs per call; more runs
 * fold pairwise on the Scala side. */
HS_JNI(jlong, mergeRuns)(JNIEnv* env, jclass cls, jlong ctx, jobject colsA, jlongArray offA, jobject colsB,
                         jlongArray offB, jint ncols, jint nIndexed, jint numBuckets) {
  (void)cls;
  view_hold a, b;
  memset(&a, 0, sizeof a);
  memset(&b, 0, sizeof b);
  hs_table* out = NULL;
  if (view_open(env, &a, colsA, ncols, nIndexed, numBuckets, offA) &&
      view_open(env, &b, colsB, ncols, nIndexed, numBuckets, offB)) {
    hs_index_view runs[2];
    runs[0] = a.v;
    runs[1] = b.v;
    int32_t st = hs_merge_runs(as_ctx(ctx), runs, 2, &out);
    if (st) { throw_hs(env, as_ctx(ctx), st); out = NULL; }
  }
  view_close(env, &a);
  view_close(env, &b);
  return (jlong)(intptr_t)out;
}

/* ---- the plan of JoinIndexRule.apply (HS/index/rules/JoinIndexRule.scala:57-90) over two index relations
 * (HS/index/rules/RuleUtils.scala:264-292). joinType: hs_join_type. */
HS_JNI(jlong, joinBucketed)(JNIEnv* env, jclass cls, jlong ctx, jobject lcols, jint lncols, jlongArray loff,
                            jobject rcols, jint rncols, jlongArray roff, jint nIndexed, jint lBuckets,
                            jint rBuckets, jint joinType, jintArray leftOut, jintArray rightOut) {
  (void)cls;
  view_hold l, r;
  memset(&l, 0, sizeof l);
  memset(&r, 0, sizeof r);
  hs_table* out = NULL;
  jint* lo = (*env)->GetIntArrayElements(env, leftOut, NULL);
  jint* ro = (*env)->GetIntArrayElements(env, rightOut, NULL);
  if (lo && ro && view_open(env, &l, lcols, lncols, nIndexed, lBuckets, loff) &&
      view_open(env, &r, rcols, rncols, nIndexed, rBuckets, roff)) {
    int32_t st = hs_join_bucketed(as_ctx(ctx), &l.v, &r.v, joinType, (const int32_t*)lo,
                                  (*env)->GetArrayLength(env, leftOut), (const int32_t*)ro,
                                  (*env)->GetArrayLength(env, rightOut), &out);
    if (st) { throw_hs(env, as_ctx(ctx), st); out = NULL; }
  }
  view_close(env, &l);
  view_close(env, &r);
  if (lo) (*env)->ReleaseIntArrayElements(env, leftOut, lo, JNI_ABORT);
  if (ro) (*env)->ReleaseIntArrayElements(env, rightOut, ro, JNI_ABORT);
  return (jlong)(intptr_t)out;
}

/* ---- the plan of FilterIndexRule.apply (HS/index/rules/FilterIndexRule.scala:43-85). terms: a direct
 * ByteBuffer holding nterms hs_pred_term structs in postfix order (string literals and IN lists point into
 * direct buffers the caller keeps alive for the call). */
HS_JNI(jlong, filterScan)(JNIEnv* env, jclass cls, jlong ctx, jobject cols, jint ncols, jlongArray off, jint nIndexed,
                          jint numBuckets, jobject terms, jint nterms, jintArray outCols) {
  (void)cls;
  view_hold v;
  memset(&v, 0, sizeof v);
  hs_table* out = NULL;
  jint* oc = (*env)->GetIntArrayElements(env, outCols, NULL);
  if (oc && view_open(env, &v, cols, ncols, nIndexed, numBuckets, off)) {
    hs_predicate pred;
    pred.nterms = nterms;
    pred.reserved = 0;
    pred.terms = (const hs_pred_term*)buf_addr(env, terms);
    int32_t st = hs_filter_scan(as_ctx(ctx), &v.v, &pred, (const int32_t*)oc, (*env)->GetArrayLength(env, outCols), &out);
    if (st) { throw_hs(env, as_ctx(ctx), st); out = NULL; }
  }
  view_close(env, &v);
  if (oc) (*env)->ReleaseIntArrayElements(env, outCols, oc, JNI_ABORT);
  return (jlong)(intptr_t)out;
}

/* ---- K4 on the device: the bodies of all Parquet data pages (and dictionary pages) of ONE column of the sorted
 * result, for all bucket files at once — what Spark's ParquetFileFormat writer does to every value behind
 * saveWithBuckets (HS/index/DataFrameWriterExtensions.scala:58-65). col: a direct ByteBuffer holding one
 * hs_column; pageRows: npages + 1 row boundaries; pageChunk: the bucket file of every page. Returns a table:
 * column 0 the bodies (pages, then one dictionary page per chunk), column 1 NULLs per page / dictionary entries
 * per chunk, column 2 (codec 1 = snappy only) the uncompressed sizes. GpuPageFileWriter.scala frames them. */
HS_JNI(jlong, parquetEncodePages)(JNIEnv* env, jclass cls, jlong ctx, jobject col, jboolean optional, jlongArray pageRows,
                                  jintArray pageChunk, jint nchunks, jint maxDict, jint codec) {
  (void)cls;
  hs_table* out = NULL;
  jlong* pr = (*env)->GetLongArrayElements(env, pageRows, NULL);
  jint* pc = (*env)->GetIntArrayElements(env, pageChunk, NULL);
  if (pr && pc) {
    const jint npages = (*env)->GetArrayLength(env, pageChunk);
    int32_t st = hs_parquet_encode_pages(as_ctx(ctx), (const hs_column*)buf_addr(env, col), optional ? 1 : 0,
                                         (const int64_t*)pr, npages, (const int32_t*)pc, nchunks, maxDict, codec, &out);
    if (st) { throw_hs(env, as_ctx(ctx), st); out = NULL; }
  }
  if (pc) (*env)->ReleaseIntArrayElements(env, pageChunk, pc, JNI_ABORT);
  if (pr) (*env)->ReleaseLongArrayElements(env, pageRows, pr, JNI_ABORT);
  return (jlong)(intptr_t)out;
}

/* ---- results: the Scala side wraps the table's buffers as direct ByteBuffers (no copy) and frees the table
 * when the Parquet writer / the Spark ColumnarBatch is done with them */
HS_JNI(void, tableWait)(JNIEnv* env, jclass cls, jlong ctx, jlong table) {
  (void)cls;
  int32_t st = hs_table_wait(as_ctx(ctx), as_table(table));
  if (st) throw_hs(env, as_ctx(ctx), st);
}
HS_JNI(void, tableFree)(JNIEnv* env, jclass cls, jlong ctx, jlong table) {
  (void)env; (void)cls;
  hs_table_free(as_ctx(ctx), as_table(table));
}
HS_JNI(jlong, tableRows)(JNIEnv* env, jclass cls, jlong table) {
  (void)env; (void)cls;
  return (jlong)as_table(table)->nrows;
}
HS_JNI(jint, tableColumns)(JNIEnv* env, jclass cls, jlong table) {
  (void)env; (void)cls;
  return as_table(table)->ncols;
}
HS_JNI(jlongArray, tableBucketOffsets)(JNIEnv* env, jclass cls, jlong table) {
  (void)cls;
  const hs_table* t = as_table(table);
  if (!t->bucket_offsets) return NULL;
  jlongArray a = (*env)->NewLongArray(env, t->num_buckets + 1);
  if (a) (*env)->SetLongArrayRegion(env, a, 0, t->num_buckets + 1, (const jlong*)t->bucket_offsets);
  return a;
}
/* {type, length, null_count, validity address, offsets address, data address} of column i */
HS_JNI(jlongArray, tableColumn)(JNIEnv* env, jclass cls, jlong table, jint i) {
  (void)cls;
  const hs_table* t = as_table(table);
  if (i < 0 || i >= t->ncols) return NULL;
  const hs_column* c = &t->cols[i];
  jlong v[6];
  v[0] = c->type;
  v[1] = c->length;
  v[2] = c->null_count;
  v[3] = (jlong)(intptr_t)c->validity;
  v[4] = (jlong)(intptr_t)c->offsets;
  v[5] = (jlong)(intptr_t)c->data;
  jlongArray a = (*env)->NewLongArray(env, 6);
  if (a) (*env)->SetLongArrayRegion(env, a, 0, 6, v);
  return a;
}
/* a host buffer of a result column as a direct ByteBuffer (valid until tableFree) */
HS_JNI(jobject, wrap)(JNIEnv* env, jclass cls, jlong address, jlong bytes) {
  (void)cls;
  return (*env)->NewDirectByteBuffer(env, (void*)(intptr_t)address, bytes);
}
