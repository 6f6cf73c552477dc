package org.uu.nl.embedding.nativepath;

import org.uu.nl.embedding.convert.util.InEdgeNeighborhoodAlgorithm;
import org.uu.nl.embedding.convert.util.OutEdgeNeighborhoodAlgorithm;
import org.uu.nl.embedding.util.CoOccurrenceMatrix;
import org.uu.nl.embedding.util.InMemoryRdfGraph;
import org.uu.nl.embedding.util.config.Configuration;

import java.util.Arrays;

/**
 * Drop-in for new BookmarkColoring(graph, config) (bca/BookmarkColoring.java:32-120): same
 * constructor arguments, same CoOccurrenceMatrix answers, the matrix itself lives in HBM.
 * getKey/getType stay on the Java side (they read the graph's label / type properties,
 * BookmarkColoring.java:139-141,158-160).
 */
public class NativeCoOccurrence implements CoOccurrenceMatrix, AutoCloseable {

    final long ctx;
    private final InMemoryRdfGraph graph;
    private final int vocabSize;
    private final long count;
    private final double max;
    private int[] I, J;        // lazily exported: only cIdx_* callers (not the native trainer) need them
    private float[] X;

    public NativeCoOccurrence(final InMemoryRdfGraph graph, final Configuration config) {
        this.graph = graph;
        this.vocabSize = graph.getVertices().toIntArray().length;

        int opt;
        switch (config.getOpt().getMethodEnum()) {
            case ADAGRAD: opt = GraphEmb.OPT_ADAGRAD; break;
            case ADAM:    opt = GraphEmb.OPT_ADAM; break;
            case AMSGRAD: opt = GraphEmb.OPT_AMSGRAD; break;
            default: throw new IllegalArgumentException("Invalid optimization method");
        }
        int method;
        switch (config.getMethodEnum()) {
            case GLOVE:  method = GraphEmb.METHOD_GLOVE; break;
            case PGLOVE: method = GraphEmb.METHOD_PGLOVE; break;
            default: throw new IllegalArgumentException("Invalid cost function");
        }
        int norm;
        switch (config.getBca().getNormalizeEnum()) {
            case UNITY:  norm = GraphEmb.NORM_UNITY; break;
            case COUNTS: norm = GraphEmb.NORM_COUNTS; break;
            default:     norm = GraphEmb.NORM_NONE;
        }
        this.ctx = GraphEmb.create(config.getDim(), method, opt, config.getBca().isDirected(), norm,
                config.getBca().getAlpha(), config.getBca().getEpsilon(),
                config.getOpt().getTolerance(), config.getOpt().getMaxiter(), System.currentTimeMillis());

        // exactly the four arrays the reference hands its jobs (BookmarkColoring.java:49-52)
        final int[][] inVertex = graph.getInNeighborhoods();
        final int[][] outVertex = graph.getOutNeighborhoods();
        final int[][] inEdge = new InEdgeNeighborhoodAlgorithm(config).compute(graph);
        final int[][] outEdge = new OutEdgeNeighborhoodAlgorithm(config).compute(graph);
        final Csr out = flatten(outVertex, outEdge), in = flatten(inVertex, inEdge);
        GraphEmb.setGraphCsr(ctx, vocabSize, out.off, out.nbr, out.w, in.off, in.nbr, in.w);

        this.count = GraphEmb.coocBuild(ctx);
        this.max = GraphEmb.coocMax(ctx);
    }

    private static final class Csr { long[] off; int[] nbr; float[] w; }

    /** jagged neighbour/edge arrays -> CSR with each list sorted by neighbour id */
    private Csr flatten(int[][] vertex, int[][] edge) {
        final Csr c = new Csr();
        c.off = new long[vocabSize + 1];
        for (int v = 0; v < vocabSize; v++) c.off[v + 1] = c.off[v] + vertex[v].length;
        final int e = (int) c.off[vocabSize];
        c.nbr = new int[e];
        c.w = new float[e];
        for (int v = 0; v < vocabSize; v++) {
            final int n = vertex[v].length;
            final long[] keyed = new long[n];
            for (int k = 0; k < n; k++) keyed[k] = ((long) vertex[v][k] << 32) | (k & 0xffffffffL);
            Arrays.sort(keyed);
            for (int k = 0; k < n; k++) {
                final int src = (int) keyed[k];
                c.nbr[(int) c.off[v] + k] = (int) (keyed[k] >>> 32);
                c.w[(int) c.off[v] + k] = graph.getEdgeWeightProperty().getValueAsFloat(edge[v][src]);
            }
        }
        return c;
    }

    private void export() {
        if (I != null) return;
        if (count > Integer.MAX_VALUE) throw new IllegalStateException("matrix too large for int-indexed access");
        I = new int[(int) count]; J = new int[(int) count]; X = new float[(int) count];
        GraphEmb.coocExport(ctx, I, J, X);
    }

    @Override public int vocabSize() { return vocabSize; }
    @Override public double max() { return max; }
    @Override public int coOccurrenceCount() { return (int) count; }
    @Override public int cIdx_I(int i) { export(); return I[i]; }
    @Override public int cIdx_J(int j) { export(); return J[j]; }
    @Override public float cIdx_C(int i) { export(); return X[i]; }
    @Override public void shuffle() { /* the native trainer owns the epoch order */ }
    @Override public byte getType(int index) { return (byte) graph.getVertexTypeProperty().getValueAsInt(index); }
    @Override public String getKey(int index) { return graph.getVertexLabelProperty().getValueAsString(index); }
    @Override public void close() { GraphEmb.destroy(ctx); }
}
