/*
 * JNI shim for hex.tree.gpu.H2OTreeB200: the C side of the native methods, one thin wrapper per
 * entry point of include/h2ot.h. Shown for four of them -- the others follow the same three lines (pin the Java array or take
 * the direct buffer's address, call h2ot_*, turn a non-zero status into a RuntimeException carrying h2ot_last_error()).
 *
 * Not built by this repository: the image has no JDK (no jni.h). With one: 
 *   gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -Iinclude integration/h2otree_jni.c \
 *       -Lh2o-3_b200 -lh2otree_b200 -o libh2otree_jni.so
 */
#include <jni.h>
#include <stdint.h>

#include "h2ot.h"

static void throw_if(JNIEnv* env, int status) {
  if (status == H2OT_OK) return;
  jclass ex = (*env)->FindClass(env, "java/lang/RuntimeException");
  if (ex) (*env)->ThrowNew(env, ex, h2ot_last_error());
}

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

JNIEXPORT void JNICALL Java_hex_tree_gpu_H2OTreeB200_destroy(JNIEnv* env, jclass cls, jlong h) {
  throw_if(env, h2ot_destroy((h2ot_engine*)(intptr_t)h));
}

/* Chunk._mem goes down as it is: no getDoubles on the JVM side. */
JNIEXPORT void JNICALL Java_hex_tree_gpu_H2OTreeB200_loadChunk(JNIEnv* env, jclass cls, jlong h, jint col, jlong row0, jlong n, jint codec,
                                                               jbyteArray mem) {
  const jsize len = (*env)->GetArrayLength(env, mem);
  jbyte* p = (jbyte*)(*env)->GetPrimitiveArrayCritical(env, mem, NULL);
  if (!p) return; /* OutOfMemoryError already pending */
  const int rc = h2ot_load_chunk((h2ot_engine*)(intptr_t)h, col, row0, n, codec, (const uint8_t*)p, len);
  (*env)->ReleasePrimitiveArrayCritical(env, mem, p, JNI_ABORT);
  throw_if(env, rc);
}

/* One tree level: decided nodes, bitsets, descriptors and the split records travel in direct ByteBuffers laid out as the C structs. */
JNIEXPORT void JNICALL Java_hex_tree_gpu_H2OTreeB200_levelStep(JNIEnv* env, jclass cls, jlong h, jint k, jint leafOffset, jint nLeaves,
                                                               jint nDecided, jobject decided, jobject bitsets, jlong bitsetBytes,
                                                               jobject descs, jdouble minRows, jdouble minSplitImprovement,
                                                               jdoubleArray min2, jdoubleArray maxIn, jdoubleArray wNA, jbyteArray hasVals,
                                                               jobject outSplits) {
  const h2ot_decided_node* dn = decided ? (const h2ot_decided_node*)(*env)->GetDirectBufferAddress(env, decided) : NULL;
  const uint8_t* bs = bitsets ? (const uint8_t*)(*env)->GetDirectBufferAddress(env, bitsets) : NULL;
  const h2ot_hist_desc* ds = (const h2ot_hist_desc*)(*env)->GetDirectBufferAddress(env, descs);
  h2ot_split* sp = (h2ot_split*)(*env)->GetDirectBufferAddress(env, outSplits);
  jdouble* m2 = (*env)->GetDoubleArrayElements(env, min2, NULL);
  jdouble* mi = (*env)->GetDoubleArrayElements(env, maxIn, NULL);
  jdouble* wn = (*env)->GetDoubleArrayElements(env, wNA, NULL);
  jbyte* hv = (*env)->GetByteArrayElements(env, hasVals, NULL);
  const int rc = h2ot_level_step((h2ot_engine*)(intptr_t)h, k, leafOffset, nLeaves, nDecided, dn, bs, bitsetBytes, ds, minRows,
                                 minSplitImprovement, m2, mi, wn, (uint8_t*)hv, sp);
  (*env)->ReleaseDoubleArrayElements(env, min2, m2, 0);
  (*env)->ReleaseDoubleArrayElements(env, maxIn, mi, 0);
  (*env)->ReleaseDoubleArrayElements(env, wNA, wn, 0);
  (*env)->ReleaseByteArrayElements(env, hasVals, hv, 0);
  throw_if(env, rc);
}
