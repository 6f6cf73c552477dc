package org.uu.nl.embedding.nativepath;

import org.uu.nl.embedding.opt.IOptimizer;
import org.uu.nl.embedding.opt.OptimizeJob;
import org.uu.nl.embedding.opt.Optimum;
import org.uu.nl.embedding.util.config.Configuration;

/**
 * Drop-in for Main.createOptimizer(config, coMatrix) (Main.java:107-131): optimize() runs the
 * whole epoch loop of opt/Optimizer.java:67-120 on the GPU and fills an Optimum exactly as the
 * reference does (result = (focus+context)/2 as double[V*dim], cost history, finalCost — which
 * stays 0 when maxiter is exhausted, Optimizer.java:73,104-107).
 */
public class NativeOptimizer implements IOptimizer {

    private final NativeCoOccurrence matrix;
    private final Configuration config;

    public NativeOptimizer(NativeCoOccurrence matrix, Configuration config) {
        this.matrix = matrix;
        this.config = config;
    }

    @Override
    public Optimum optimize() {
        final Optimum opt = new Optimum();
        final double[] r = GraphEmb.train(matrix.ctx, (epoch, cost, seconds) -> false);
        for (double c : GraphEmb.costHistory(matrix.ctx)) opt.addIntermediaryResult(c);
        opt.setResult(extractResult());
        opt.setFinalCost(r[0]);
        return opt;
    }

    @Override public String getName() { return config.getOpt().getMethod() + " (native, sm_100a)"; }

    /** the per-thread job objects of the reference do not exist on the device */
    @Override public OptimizeJob createJob(int id, int iteration) {
        throw new UnsupportedOperationException("the native trainer runs whole epochs (ge_train_epoch)");
    }

    @Override
    public double[] extractResult() {
        final double[] out = new double[matrix.vocabSize() * config.getDim()];
        GraphEmb.extract(matrix.ctx, out);
        return out;
    }
}
