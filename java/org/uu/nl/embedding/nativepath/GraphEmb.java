package org.uu.nl.embedding.nativepath;

/**
 * JNI declarations for libgraphemb.so (include/graphemb.h).  One static native per C entry
 * point; the C side is java/jni/graphemb_jni.c.  Targets Java 8 (pom.xml:13-14), so JNI rather
 * than Panama FFM; INTEGRATION.md shows the FFM downcall handles for JDK >= 22.
 *
 * NOT COMPILED IN THIS REPOSITORY'S CI: the build image has no JDK.  The exercised callers of
 * the C ABI are tests/ (Python ctypes) — this file is the binding a maintainer of the
 * reference adds.
 */
public final class GraphEmb {

    static { System.loadLibrary("graphemb_jni"); }   // links against libgraphemb.so

    private GraphEmb() {}

    public static final int METHOD_GLOVE = 0, METHOD_PGLOVE = 1;
    public static final int OPT_ADAGRAD = 0, OPT_ADAM = 1, OPT_AMSGRAD = 2;
    public static final int NORM_NONE = 0, NORM_UNITY = 1, NORM_COUNTS = 2;

    /** ge_create: returns the opaque handle, throws IllegalStateException with ge_last_error. */
    public static native long create(int dim, int method, int optMethod, boolean directed, int normalize,
                                     double alpha, double epsilon, double tolerance, int maxiter, long seed);
    public static native void destroy(long ctx);

    /** ge_set_graph_csr: neighbour lists sorted by id, unique. */
    public static native void setGraphCsr(long ctx, int vocab, long[] outOff, int[] outNbr, float[] outW,
                                          long[] inOff, int[] inNbr, float[] inW);
    /** ge_set_graph_edges: raw directed edge list, flattened on the device. */
    public static native void setGraphEdges(long ctx, int vocab, int[] src, int[] dst, float[] w);

    /** ge_cooc_build: returns coOccurrenceCount(); max() via coocMax. */
    public static native long coocBuild(long ctx);
    public static native double coocMax(long ctx);
    public static native void coocExport(long ctx, int[] i, int[] j, float[] x);

    /** ge_train with an upcall per epoch (may be null). Returns {finalCost, epochs}. */
    public interface EpochListener { boolean onEpoch(int epoch, double cost, double seconds); }
    public static native double[] train(long ctx, EpochListener listener);
    public static native double[] costHistory(long ctx);
    /** ge_extract: (focus+context)/2 as double[V*dim], row-major (Optimizer.extractResult). */
    public static native void extract(long ctx, double[] out);

    /** ge_similarity_pairs + ge_similarity_export: the CompareJob loop of one CompareGroup on the device.
     *  Arrays are the fields of ge_sim_group (include/graphemb.h); unused ones may be null.
     *  Returns {vert[], other[]} as int[2][count]; the similarities are written to sim[0]. */
    public static native int[][] similarityPairs(long ctx, int method, boolean upperTriangle, double threshold,
                                                 double smooth, double distance, int timeDir, int[] vert, int[] stringId,
                                                 byte[] flags, long[] profOff, int[] profKey, int[] profCnt, int nCut,
                                                 int[] cutSlot, int[] numVal, byte[] numStatus, long[] epochDay,
                                                 long[] packedMonth, int[] source, int[] target, long[] charOff, char[] chars,
                                                 float[][] sim);

    /** ge_write_embeddings: EmbeddingTextWriter.write from the device store. */
    public static native void writeEmbeddings(long ctx, String vectorsPath, String dictPath, String header,
                                              String[] keys, byte[] types, boolean[] typeEnabled,
                                              String[][] prefixes);
}
