/*
 * JNI shim for hex.tree.gpu.H2OTreeB200: the C side of every native method, one thin wrapper per entry
 * point of include/h2ot.h. Each wrapper pins the Java arrays (or takes the direct buffers' addresses), calls h2ot_*, and turns a
 * non-zero status into a RuntimeException that carries h2ot_last_error().
 *
 * Not built by this repository: the image has no JDK (no jni.h), so this file has only been syntax-checked against a stand-in
 * header with the JNI types it uses (tests/test_capi_load.py::test_jni_shim_compiles_against_stub_header). With a JDK:
 *   gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -Iinclude integration/h2otree_jni.c \
 *       -Lh2o-3_b200 -lh2otree_b200 -o libh2otree_jni.so
 * Array arguments: bulk row data (column values, predictions, nids) is pinned with GetPrimitiveArrayCritical only around calls
 * that do not block on the JVM; Chunk bytes are copied out with Get<Type>ArrayElements because h2ot_load_chunk synchronises
 * with the device. Records (decided nodes, descriptors, splits) travel in direct ByteBuffers laid out as the C structs.
 */
#include <jni.h>
#include <stdint.h>
#include <stddef.h>

#include "h2ot.h"

#define H(h) ((h2ot_engine*)(intptr_t)(h))
#define FN(name) JNIEXPORT void JNICALL Java_hex_tree_gpu_H2OTreeB200_##name

static void throw_if(JNIEnv* env, int status) {
  if (status == H2OT_OK) return;
  jclass ex = (*env)->FindClass(env, "java/lang/RuntimeException");
  if (ex) (*env)->ThrowNew(env, ex, h2ot_last_error());
}
static void* buf(JNIEnv* env, jobject b) { return b ? (*env)->GetDirectBufferAddress(env, b) : NULL; }
/* Get<Type>ArrayElements with the NULL check every caller needs (an OutOfMemoryError is then pending) */
#define PIN_D(var, arr) jdouble* var = (arr) ? (*env)->GetDoubleArrayElements(env, (arr), NULL) : NULL; if ((arr) && !var) return
#define UNPIN_D(var, arr, mode) if (var) (*env)->ReleaseDoubleArrayElements(env, (arr), var, (mode))
#define PIN_F(var, arr) jfloat* var = (arr) ? (*env)->GetFloatArrayElements(env, (arr), NULL) : NULL; if ((arr) && !var) return
#define UNPIN_F(var, arr, mode) if (var) (*env)->ReleaseFloatArrayElements(env, (arr), var, (mode))
#define PIN_B(var, arr) jbyte* var = (arr) ? (*env)->GetByteArrayElements(env, (arr), NULL) : NULL; if ((arr) && !var) return
#define UNPIN_B(var, arr, mode) if (var) (*env)->ReleaseByteArrayElements(env, (arr), var, (mode))
#define PIN_I(var, arr) jint* var = (arr) ? (*env)->GetIntArrayElements(env, (arr), NULL) : NULL; if ((arr) && !var) return
#define UNPIN_I(var, arr, mode) if (var) (*env)->ReleaseIntArrayElements(env, (arr), var, (mode))
#define PIN_L(var, arr) jlong* var = (arr) ? (*env)->GetLongArrayElements(env, (arr), NULL) : NULL; if ((arr) && !var) return
#define UNPIN_L(var, arr, mode) if (var) (*env)->ReleaseLongArrayElements(env, (arr), var, (mode))

JNIEXPORT jlong JNICALL Java_hex_tree_gpu_H2OTreeB200_create(JNIEnv* env, jclass cls, jint device, jlong nrows, jint ncols, jlong rowOffset,
                                                             jlong nrowsGlobal, jint nclassTrees, jint nclass, jint family,
                                                             jboolean hasWeights, jboolean hasOffset) {
  h2ot_config cfg = {0};
  cfg.device = device;
  cfg.nrows = nrows;
  cfg.ncols = ncols;
  cfg.row_offset = rowOffset;
  cfg.nrows_global = nrowsGlobal;
  cfg.nclass_trees = nclassTrees;
  cfg.nclass = nclass;
  cfg.family = family;
  cfg.has_weights = hasWeights;
  cfg.has_offset = hasOffset;
  cfg.tweedie_power = 1.5;
  h2ot_engine* h = NULL;
  throw_if(env, h2ot_create(&cfg, &h));
  return (jlong)(intptr_t)h;
}

FN(destroy)(JNIEnv* env, jclass cls, jlong h) { throw_if(env, h2ot_destroy(H(h))); }

/* ---- frame ingest */
FN(loadColumn)(JNIEnv* env, jclass cls, jlong h, jint col, jlong row0, jdoubleArray vals, jint n) {
  PIN_D(v, vals);
  const int rc = h2ot_load_column(H(h), col, row0, n, v);
  UNPIN_D(v, vals, JNI_ABORT);
  throw_if(env, rc);
}
/* Chunk._mem goes down as it is: no getDoubles on the JVM side. */
FN(loadChunk)(JNIEnv* env, jclass cls, jlong h, jint col, jlong row0, jlong n, jint codec, jbyteArray mem) {
  const jsize len = (*env)->GetArrayLength(env, mem);
  PIN_B(p, mem);
  const int rc = h2ot_load_chunk(H(h), col, row0, n, codec, (const uint8_t*)p, len);
  UNPIN_B(p, mem, JNI_ABORT);
  throw_if(env, rc);
}
FN(loadResponse)(JNIEnv* env, jclass cls, jlong h, jlong row0, jdoubleArray vals, jint n) {
  PIN_D(v, vals);
  const int rc = h2ot_load_response(H(h), row0, n, v);
  UNPIN_D(v, vals, JNI_ABORT);
  throw_if(env, rc);
}
FN(loadWeights)(JNIEnv* env, jclass cls, jlong h, jlong row0, jdoubleArray vals, jint n) {
  PIN_D(v, vals);
  const int rc = h2ot_load_weights(H(h), row0, n, v);
  UNPIN_D(v, vals, JNI_ABORT);
  throw_if(env, rc);
}
FN(loadOffset)(JNIEnv* env, jclass cls, jlong h, jlong row0, jdoubleArray vals, jint n) {
  PIN_D(v, vals);
  const int rc = h2ot_load_offset(H(h), row0, n, v);
  UNPIN_D(v, vals, JNI_ABORT);
  throw_if(env, rc);
}
FN(finalizeFrame)(JNIEnv* env, jclass cls, jlong h) { throw_if(env, h2ot_finalize_frame(H(h))); }

/* ---- model state */
JNIEXPORT jlong JNICALL Java_hex_tree_gpu_H2OTreeB200_initModel(JNIEnv* env, jclass cls, jlong h, jdoubleArray initF) {
  jdouble* f = (*env)->GetDoubleArrayElements(env, initF, NULL);
  if (!f) return 0;
  int64_t skipped = 0;
  const int rc = h2ot_init_model(H(h), f, &skipped);
  (*env)->ReleaseDoubleArrayElements(env, initF, f, JNI_ABORT);
  throw_if(env, rc);
  return (jlong)skipped;
}
FN(initialStats)(JNIEnv* env, jclass cls, jlong h, jdoubleArray numDen, jdoubleArray classWeight, jdoubleArray absMax) {
  PIN_D(nd, numDen);
  PIN_D(cw, classWeight);
  PIN_D(am, absMax);
  const int rc = h2ot_initial_stats(H(h), nd, nd + 1, cw, am);
  UNPIN_D(nd, numDen, 0);
  UNPIN_D(cw, classWeight, 0);
  UNPIN_D(am, absMax, 0);
  throw_if(env, rc);
}
FN(computeResiduals)(JNIEnv* env, jclass cls, jlong h, jdoubleArray outAbsMax) {
  PIN_D(v, outAbsMax);
  const int rc = h2ot_compute_residuals(H(h), v);
  UNPIN_D(v, outAbsMax, 0);
  throw_if(env, rc);
}
FN(beginTree)(JNIEnv* env, jclass cls, jlong h, jint k, jlong treeSeed, jdouble sampleRate, jdouble workAbsBound) {
  throw_if(env, h2ot_begin_tree(H(h), k, treeSeed, sampleRate, workAbsBound));
}
FN(getPredictions)(JNIEnv* env, jclass cls, jlong h, jint k, jdoubleArray out) {
  PIN_D(v, out);
  const int rc = h2ot_get_predictions(H(h), k, v);
  UNPIN_D(v, out, 0);
  throw_if(env, rc);
}
FN(getNids)(JNIEnv* env, jclass cls, jlong h, jint k, jintArray out) {
  PIN_I(v, out);
  const int rc = h2ot_get_nids(H(h), k, (int32_t*)v);
  UNPIN_I(v, out, 0);
  throw_if(env, rc);
}

/* ---- one tree level */
FN(buildLevel)(JNIEnv* env, jclass cls, jlong h, jint k, jint leafOffset, jint nLeaves, jint nDecided, jobject decided, jobject bitsets,
               jlong bitsetBytes, jobject descs) {
  throw_if(env, h2ot_build_level(H(h), k, leafOffset, nLeaves, nDecided, (const h2ot_decided_node*)buf(env, decided),
                                 (const uint8_t*)buf(env, bitsets), bitsetBytes, (const h2ot_hist_desc*)buf(env, descs)));
}
FN(setSplitPoints)(JNIEnv* env, jclass cls, jlong h, jint k, jdoubleArray points) {
  const jsize n = (*env)->GetArrayLength(env, points);
  PIN_D(v, points);
  const int rc = h2ot_set_split_points(H(h), k, v, n);
  UNPIN_D(v, points, JNI_ABORT);
  throw_if(env, rc);
}
FN(allreduceLevel)(JNIEnv* env, jclass cls, jlong h, jint k) { throw_if(env, h2ot_allreduce_level(H(h), k)); }
FN(finishLevel)(JNIEnv* env, jclass cls, jlong h, jint k, jdoubleArray vals, jlongArray valOffsets, jdoubleArray min2, jdoubleArray maxIn,
                jdoubleArray wNA, jbyteArray hasVals) {
  const jsize cap = vals ? (*env)->GetArrayLength(env, vals) : 0;
  PIN_D(v, vals);
  PIN_L(vo, valOffsets);
  PIN_D(m2, min2);
  PIN_D(mi, maxIn);
  PIN_D(wn, wNA);
  PIN_B(hv, hasVals);
  const int rc = h2ot_finish_level(H(h), k, v, cap, (int64_t*)vo, m2, mi, wn, (uint8_t*)hv);
  UNPIN_D(v, vals, 0);
  UNPIN_L(vo, valOffsets, 0);
  UNPIN_D(m2, min2, 0);
  UNPIN_D(mi, maxIn, 0);
  UNPIN_D(wn, wNA, 0);
  UNPIN_B(hv, hasVals, 0);
  throw_if(env, rc);
}
FN(findSplits)(JNIEnv* env, jclass cls, jlong h, jint k, jdouble minRows, jdouble minSplitImprovement, jobject outSplits) {
  throw_if(env, h2ot_find_splits(H(h), k, minRows, minSplitImprovement, (h2ot_split*)buf(env, outSplits)));
}
/* buildLevel + allreduceLevel + finishLevel (no histogram download) + findSplits with one synchronisation */
FN(levelStep)(JNIEnv* env, jclass cls, jlong h, jint k, jint leafOffset, jint nLeaves, jint nDecided, jobject decided, jobject bitsets,
              jlong bitsetBytes, jobject descs, jdouble minRows, jdouble minSplitImprovement, jdoubleArray min2, jdoubleArray maxIn,
              jdoubleArray wNA, jbyteArray hasVals, jobject outSplits) {
  PIN_D(m2, min2);
  PIN_D(mi, maxIn);
  PIN_D(wn, wNA);
  PIN_B(hv, hasVals);
  const int rc = h2ot_level_step(H(h), k, leafOffset, nLeaves, nDecided, (const h2ot_decided_node*)buf(env, decided),
                                 (const uint8_t*)buf(env, bitsets), bitsetBytes, (const h2ot_hist_desc*)buf(env, descs), minRows,
                                 minSplitImprovement, m2, mi, wn, (uint8_t*)hv, (h2ot_split*)buf(env, outSplits));
  UNPIN_D(m2, min2, 0);
  UNPIN_D(mi, maxIn, 0);
  UNPIN_D(wn, wNA, 0);
  UNPIN_B(hv, hasVals, 0);
  throw_if(env, rc);
}
FN(levelStepBegin)(JNIEnv* env, jclass cls, jlong h, jint k, jint leafOffset, jint nLeaves, jint nDecided, jobject decided,
                   jobject bitsets, jlong bitsetBytes, jobject descs, jdouble minRows, jdouble minSplitImprovement) {
  throw_if(env, h2ot_level_step_begin(H(h), k, leafOffset, nLeaves, nDecided, (const h2ot_decided_node*)buf(env, decided),
                                      (const uint8_t*)buf(env, bitsets), bitsetBytes, (const h2ot_hist_desc*)buf(env, descs), minRows,
                                      minSplitImprovement));
}
FN(levelStepEnd)(JNIEnv* env, jclass cls, jlong h, jint k, jdoubleArray min2, jdoubleArray maxIn, jdoubleArray wNA, jbyteArray hasVals,
                 jobject outSplits) {
  PIN_D(m2, min2);
  PIN_D(mi, maxIn);
  PIN_D(wn, wNA);
  PIN_B(hv, hasVals);
  const int rc = h2ot_level_step_end(H(h), k, m2, mi, wn, (uint8_t*)hv, (h2ot_split*)buf(env, outSplits));
  UNPIN_D(m2, min2, 0);
  UNPIN_D(mi, maxIn, 0);
  UNPIN_D(wn, wNA, 0);
  UNPIN_B(hv, hasVals, 0);
  throw_if(env, rc);
}

/* ---- per-tree row passes */
FN(setTree)(JNIEnv* env, jclass cls, jlong h, jint k, jint nNodes, jobject nodes, jobject bitsets, jlong bitsetBytes) {
  throw_if(env, h2ot_set_tree(H(h), k, nNodes, (const h2ot_decided_node*)buf(env, nodes), (const uint8_t*)buf(env, bitsets), bitsetBytes));
}
FN(gammaPass)(JNIEnv* env, jclass cls, jlong h, jint k, jint leafStart, jint nLeafNodes, jdoubleArray num, jdoubleArray denom) {
  PIN_D(a, num);
  PIN_D(b, denom);
  const int rc = h2ot_gamma_pass(H(h), k, leafStart, nLeafNodes, a, b);
  UNPIN_D(a, num, 0);
  UNPIN_D(b, denom, 0);
  throw_if(env, rc);
}
FN(addTree)(JNIEnv* env, jclass cls, jlong h, jint k, jfloatArray nodePred) {
  const jsize n = (*env)->GetArrayLength(env, nodePred);
  PIN_F(v, nodePred);
  const int rc = h2ot_add_tree(H(h), k, n, v);
  UNPIN_F(v, nodePred, JNI_ABORT);
  throw_if(env, rc);
}
FN(scoreTree)(JNIEnv* env, jclass cls, jlong h, jint k, jint leafStart, jint nNodes, jfloatArray nodePred) {
  PIN_F(v, nodePred);
  const int rc = h2ot_score_tree(H(h), k, leafStart, nNodes, v);
  UNPIN_F(v, nodePred, JNI_ABORT);
  throw_if(env, rc);
}
FN(scoreMetrics)(JNIEnv* env, jclass cls, jlong h, jdoubleArray out4) {
  h2ot_metrics m;
  const int rc = h2ot_score_metrics(H(h), &m);
  if (rc == H2OT_OK) (*env)->SetDoubleArrayRegion(env, out4, 0, (jsize)(sizeof(m) / sizeof(double) < 4 ? sizeof(m) / sizeof(double) : 4), (const jdouble*)&m);
  throw_if(env, rc);
}

/* ---- DRF */
FN(drfInitModel)(JNIEnv* env, jclass cls, jlong h) { throw_if(env, h2ot_drf_init_model(H(h))); }
FN(sample)(JNIEnv* env, jclass cls, jlong h, jint k, jlong treeSeed, jdouble sampleRate) { throw_if(env, h2ot_sample(H(h), k, treeSeed, sampleRate)); }
FN(drfCollectPreds)(JNIEnv* env, jclass cls, jlong h, jint k, jfloatArray nodePred, jboolean rootIsLeaf, jboolean updateOobt) {
  const jsize n = (*env)->GetArrayLength(env, nodePred);
  PIN_F(v, nodePred);
  const int rc = h2ot_drf_collect_preds(H(h), k, n, v, rootIsLeaf ? 1 : 0, updateOobt ? 1 : 0);
  UNPIN_F(v, nodePred, JNI_ABORT);
  throw_if(env, rc);
}
FN(getOobt)(JNIEnv* env, jclass cls, jlong h, jdoubleArray out) {
  PIN_D(v, out);
  const int rc = h2ot_get_oobt(H(h), v);
  UNPIN_D(v, out, 0);
  throw_if(env, rc);
}

/* ---- round 2: a whole tree per call, trees in a pipeline, sharded builds */
FN(treeLevelsBegin)(JNIEnv* env, jclass cls, jlong h, jint k, jint maxDepth, jobject rootDescs, jdouble minRows, jdouble minSplitImprovement,
                    jint nbins, jint nbinsCats) {
  throw_if(env, h2ot_tree_levels_begin(H(h), k, maxDepth, (const h2ot_hist_desc*)buf(env, rootDescs), minRows, minSplitImprovement, nbins,
                                       nbinsCats));
}
/* outSplits / outExists / outLeafPred: direct buffers of 2^D - 1 h2ot_split, 2^(D+1) - 1 bytes, 2^(D+1) - 1 floats (the last nullable) */
FN(treeLevelsEnd)(JNIEnv* env, jclass cls, jlong h, jint k, jobject outSplits, jobject outExists, jobject outLeafPred) {
  throw_if(env, h2ot_tree_levels_end2(H(h), k, (h2ot_split*)buf(env, outSplits), (uint8_t*)buf(env, outExists), (float*)buf(env, outLeafPred)));
}
FN(treeFinishEnqueue)(JNIEnv* env, jclass cls, jlong h, jint k, jdouble rate, jdouble maxAbsLeafnodePred) {
  throw_if(env, h2ot_tree_finish_enqueue(H(h), k, rate, maxAbsLeafnodePred));
}
FN(computeResidualsAsync)(JNIEnv* env, jclass cls, jlong h, jboolean wantBounds) {
  throw_if(env, h2ot_compute_residuals_async(H(h), wantBounds ? 1 : 0));
}
/* outAbsMax[K] and outGammaBounds[2] (both nullable) */
FN(residualBoundsWait)(JNIEnv* env, jclass cls, jlong h, jdoubleArray outAbsMax, jdoubleArray outGammaBounds) {
  PIN_D(a, outAbsMax);
  jdouble* b = outGammaBounds ? (*env)->GetDoubleArrayElements(env, outGammaBounds, NULL) : NULL;
  const int rc = h2ot_residual_bounds_wait(H(h), a, b);
  if (b) (*env)->ReleaseDoubleArrayElements(env, outGammaBounds, b, 0);
  UNPIN_D(a, outAbsMax, 0);
  throw_if(env, rc);
}
FN(setGammaBounds)(JNIEnv* env, jclass cls, jlong h, jdouble numBound, jdouble denomBound) {
  const double b[2] = {numBound, denomBound};
  throw_if(env, h2ot_set_gamma_bounds(H(h), b));
}
FN(addTreeScaled)(JNIEnv* env, jclass cls, jlong h, jint k, jfloatArray nodePred, jdoubleArray nodeFactor) {
  const jsize n = (*env)->GetArrayLength(env, nodePred);
  PIN_F(v, nodePred);
  jdouble* f = (*env)->GetDoubleArrayElements(env, nodeFactor, NULL);
  const int rc = f ? h2ot_add_tree_scaled(H(h), k, n, v, f) : H2OT_ERR_INVALID;
  if (f) (*env)->ReleaseDoubleArrayElements(env, nodeFactor, f, JNI_ABORT);
  UNPIN_F(v, nodePred, JNI_ABORT);
  throw_if(env, rc);
}
FN(setSampleRatePerClass)(JNIEnv* env, jclass cls, jlong h, jdoubleArray rates) {
  const jsize n = rates ? (*env)->GetArrayLength(env, rates) : 0;
  PIN_D(v, rates);
  const int rc = h2ot_set_sample_rate_per_class(H(h), v, n);
  UNPIN_D(v, rates, JNI_ABORT);
  throw_if(env, rc);
}
/* out: direct buffer of maxChunks * H2OT_AUC_STRIDE doubles; returns {nChunks, chunkRows} through outInfo[2] */
FN(aucBuilders)(JNIEnv* env, jclass cls, jlong h, jint maxChunks, jlongArray outInfo, jobject out) {
  int32_t n_chunks = 0;
  int64_t chunk_rows = 0;
  const int rc = h2ot_auc_builders(H(h), maxChunks, &n_chunks, &chunk_rows, (double*)buf(env, out));
  if (rc == H2OT_OK && outInfo) {
    const jlong info[2] = {n_chunks, chunk_rows};
    (*env)->SetLongArrayRegion(env, outInfo, 0, 2, info);
  }
  throw_if(env, rc);
}
FN(setLazySync)(JNIEnv* env, jclass cls, jlong h, jboolean on) { throw_if(env, h2ot_set_lazy_sync(H(h), on ? 1 : 0)); }
FN(setTimers)(JNIEnv* env, jclass cls, jlong h, jint mode) { throw_if(env, h2ot_set_timers(H(h), mode)); }
FN(reserveExchange)(JNIEnv* env, jclass cls, jlong h, jlong maxLevelBins, jlong maxLevelPairs, jlong maxLeaves) {
  throw_if(env, h2ot_reserve_exchange(H(h), maxLevelBins, maxLevelPairs, maxLeaves));
}
/* id: byte[H2OT_COMM_ID_BYTES], made on one node with commUniqueId and sent to the others (e.g. through the DKV) */
FN(commUniqueId)(JNIEnv* env, jclass cls, jbyteArray outId) {
  uint8_t id[H2OT_COMM_ID_BYTES];
  const int rc = h2ot_comm_unique_id(id);
  if (rc == H2OT_OK) (*env)->SetByteArrayRegion(env, outId, 0, H2OT_COMM_ID_BYTES, (const jbyte*)id);
  throw_if(env, rc);
}
FN(commInit)(JNIEnv* env, jclass cls, jlong h, jint worldSize, jint rank, jbyteArray id) {
  PIN_B(v, id);
  const int rc = h2ot_comm_init(H(h), worldSize, rank, (const uint8_t*)v);
  UNPIN_B(v, id, JNI_ABORT);
  throw_if(env, rc);
}
/* laplace / quantile GBM: per-leaf quantiles of the residuals (and, with useTree == false and one stratum, the initial prediction) */
FN(leafQuantiles)(JNIEnv* env, jclass cls, jlong h, jint k, jint leafStart, jint nStrata, jboolean useTree, jdouble prob, jdoubleArray out) {
  PIN_D(v, out);
  const int rc = h2ot_leaf_quantiles(H(h), k, leafStart, nStrata, useTree ? 1 : 0, prob, v);
  UNPIN_D(v, out, 0);
  throw_if(env, rc);
}
