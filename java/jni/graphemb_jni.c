/* JNI glue between org.uu.nl.embedding.nativepath.GraphEmb and libgraphemb.so.
 * Build (on a machine with a JDK):
 *   gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -I../../include \
 *       graphemb_jni.c -L../../graph_embeddings_b200 -lgraphemb -o libgraphemb_jni.so
 * Not built in this repository (no JDK in the image); shown for the maintainer. */
#include <jni.h>
#include <stdlib.h>
#include "graphemb.h"

static void fail(JNIEnv* env, ge_ctx* ctx) {
    jclass ex = (*env)->FindClass(env, "java/lang/IllegalStateException");
    (*env)->ThrowNew(env, ex, ge_last_error(ctx));
}
#define CTX ((ge_ctx*)(intptr_t)h)

JNIEXPORT jlong JNICALL Java_org_uu_nl_embedding_nativepath_GraphEmb_create(
    JNIEnv* env, jclass c, jint dim, jint method, jint opt, jboolean directed, jint norm, jdouble alpha,
    jdouble eps, jdouble tol, jint maxiter, jlong seed) {
    ge_config cfg;
    ge_config_default(&cfg);
    cfg.dim = dim; cfg.method = method; cfg.opt_method = opt; cfg.directed = directed; cfg.normalize = norm;
    cfg.alpha = alpha; cfg.epsilon = eps; cfg.tolerance = tol; cfg.maxiter = maxiter; cfg.seed = seed;
    ge_ctx* ctx = NULL;
    if (ge_create(&cfg, &ctx) != GE_OK) { fail(env, NULL); return 0; }
    return (jlong)(intptr_t)ctx;
}

JNIEXPORT void JNICALL Java_org_uu_nl_embedding_nativepath_GraphEmb_destroy(JNIEnv* env, jclass c, jlong h) {
    ge_destroy(CTX);
}

JNIEXPORT void JNICALL Java_org_uu_nl_embedding_nativepath_GraphEmb_setGraphCsr(
    JNIEnv* env, jclass c, jlong h, jint V, jlongArray oo, jintArray on, jfloatArray ow, jlongArray io,
    jintArray in, jfloatArray iw) {
    /* critical sections: the library copies to the device before returning and keeps no pointer */
    jlong* poo = (*env)->GetPrimitiveArrayCritical(env, oo, 0);
    jint* pon = (*env)->GetPrimitiveArrayCritical(env, on, 0);
    jfloat* pow_ = (*env)->GetPrimitiveArrayCritical(env, ow, 0);
    jlong* pio = (*env)->GetPrimitiveArrayCritical(env, io, 0);
    jint* pin = (*env)->GetPrimitiveArrayCritical(env, in, 0);
    jfloat* piw = (*env)->GetPrimitiveArrayCritical(env, iw, 0);
    int rc = ge_set_graph_csr(CTX, V, (const int64_t*)poo, (const int32_t*)pon, pow_, (const int64_t*)pio,
                              (const int32_t*)pin, piw);
    (*env)->ReleasePrimitiveArrayCritical(env, iw, piw, JNI_ABORT);
    (*env)->ReleasePrimitiveArrayCritical(env, in, pin, JNI_ABORT);
    (*env)->ReleasePrimitiveArrayCritical(env, io, pio, JNI_ABORT);
    (*env)->ReleasePrimitiveArrayCritical(env, ow, pow_, JNI_ABORT);
    (*env)->ReleasePrimitiveArrayCritical(env, on, pon, JNI_ABORT);
    (*env)->ReleasePrimitiveArrayCritical(env, oo, poo, JNI_ABORT);
    if (rc) fail(env, CTX);
}

JNIEXPORT jlong JNICALL Java_org_uu_nl_embedding_nativepath_GraphEmb_coocBuild(JNIEnv* env, jclass c, jlong h) {
    int64_t nnz = 0; double xmax = 0;
    if (ge_cooc_build(CTX, &nnz, &xmax)) { fail(env, CTX); return 0; }
    return nnz;
}

JNIEXPORT jdouble JNICALL Java_org_uu_nl_embedding_nativepath_GraphEmb_coocMax(JNIEnv* env, jclass c, jlong h) {
    return ge_cooc_max(CTX);
}

struct upcall { JNIEnv* env; jobject listener; jmethodID mid; };
static int epoch_cb(void* user, int32_t epoch, double cost, double seconds) {
    struct upcall* u = user;   /* ge_train calls back on the calling thread: env is valid */
    if (!u->listener) return 0;
    return (*u->env)->CallBooleanMethod(u->env, u->listener, u->mid, (jint)epoch, cost, seconds) ? 1 : 0;
}

JNIEXPORT jdoubleArray JNICALL Java_org_uu_nl_embedding_nativepath_GraphEmb_train(JNIEnv* env, jclass c, jlong h,
                                                                                jobject listener) {
    struct upcall u = {env, listener, NULL};
    if (listener) u.mid = (*env)->GetMethodID(env, (*env)->GetObjectClass(env, listener), "onEpoch", "(IDD)Z");
    double fc = 0; int32_t epochs = 0;
    if (ge_train(CTX, epoch_cb, &u, &fc, &epochs)) { fail(env, CTX); return NULL; }
    jdouble r[2] = {fc, (double)epochs};
    jdoubleArray out = (*env)->NewDoubleArray(env, 2);
    (*env)->SetDoubleArrayRegion(env, out, 0, 2, r);
    return out;
}

JNIEXPORT jdoubleArray JNICALL Java_org_uu_nl_embedding_nativepath_GraphEmb_costHistory(JNIEnv* env, jclass c, jlong h) {
    int n = ge_cost_history(CTX, NULL, 0);
    double* buf = malloc(sizeof(double) * (n > 0 ? n : 1));
    ge_cost_history(CTX, buf, n);
    jdoubleArray out = (*env)->NewDoubleArray(env, n);
    (*env)->SetDoubleArrayRegion(env, out, 0, n, buf);
    free(buf);
    return out;
}

JNIEXPORT void JNICALL Java_org_uu_nl_embedding_nativepath_GraphEmb_extract(JNIEnv* env, jclass c, jlong h,
                                                                          jdoubleArray out) {
    jdouble* p = (*env)->GetPrimitiveArrayCritical(env, out, 0);
    int rc = ge_extract(CTX, p);
    (*env)->ReleasePrimitiveArrayCritical(env, out, p, 0);
    if (rc) fail(env, CTX);
}
/* setGraphEdges, coocExport and writeEmbeddings follow the same pattern (omitted for brevity:
 * pin arrays, call ge_set_graph_edges / ge_cooc_export / ge_write_embeddings, release). */
