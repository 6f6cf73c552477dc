/* abi_driver.c — a plain-C caller of libgraphemb.so, the way a JNI glue library or any other
 * foreign host would use it: include/graphemb.h only, caller-owned buffers, int status returns.
 *
 *   abi_driver <graph.bin> <out.bin> <dim> <epochs> <seed>
 *
 * graph.bin : int32 V, int64 E_out, int64 E_in, then out_off[V+1] (i64), out_nbr (i32), out_w (f32),
 *             in_off[V+1], in_nbr, in_w                      (the CSR hand-off of ge_set_graph_csr)
 * out.bin   : int64 nnz, double xmax, I[nnz] (i32), J[nnz] (i32), X[nnz] (f32), int32 epochs_run,
 *             double final_cost, history[epochs_run] (f64), result[V*dim] (f64)
 *
 * Runs the reference's single-thread trajectory (math = java, order = java, init = java, serial).
 * Exit codes: 0 ok; 3 = ge_create reported GE_ERR_NODEVICE (no GPU: there is no CPU fallback);
 * 1 = any other failure (message on stderr).  Built and run by tests/test_abi_c_driver.py. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "graphemb.h"

static int epochs_seen = 0;
static int on_epoch(void* user, int32_t epoch, double cost, double seconds) {
    (void)user; (void)cost; (void)seconds;
    epochs_seen = epoch + 1;
    return 0; /* do not stop early */
}

#define CK(call)                                                                    \
    do {                                                                            \
        int rc__ = (call);                                                          \
        if (rc__ != GE_OK) {                                                        \
            fprintf(stderr, "%s -> %d: %s\n", #call, rc__, ge_last_error(ctx));     \
            return 1;                                                               \
        }                                                                           \
    } while (0)

static void* slurp(FILE* f, size_t bytes) {
    void* p = malloc(bytes ? bytes : 1);
    if (!p || fread(p, 1, bytes, f) != bytes) { fprintf(stderr, "short read\n"); exit(1); }
    return p;
}

int main(int argc, char** argv) {
    if (argc != 6) { fprintf(stderr, "usage: abi_driver graph.bin out.bin dim epochs seed\n"); return 1; }
    ge_config cfg;
    ge_config_default(&cfg);
    cfg.dim = atoi(argv[3]);
    cfg.method = GE_METHOD_GLOVE;
    cfg.maxiter = atoi(argv[4]);
    cfg.tolerance = 0.0;
    cfg.seed = atoll(argv[5]);
    cfg.math = GE_MATH_JAVA;
    cfg.order = GE_ORDER_JAVA;
    cfg.init = GE_INIT_JAVA;
    cfg.serial = 1;
    ge_ctx* ctx = NULL;
    int rc = ge_create(&cfg, &ctx);
    if (rc == GE_ERR_NODEVICE) { fprintf(stderr, "no device: %s\n", ge_last_error(NULL)); return 3; }
    if (rc != GE_OK) { fprintf(stderr, "ge_create -> %d: %s\n", rc, ge_last_error(NULL)); return 1; }

    FILE* f = fopen(argv[1], "rb");
    if (!f) { perror(argv[1]); return 1; }
    int32_t V; int64_t Eo, Ei;
    if (fread(&V, 4, 1, f) != 1 || fread(&Eo, 8, 1, f) != 1 || fread(&Ei, 8, 1, f) != 1) return 1;
    int64_t* out_off = slurp(f, (size_t)(V + 1) * 8);
    int32_t* out_nbr = slurp(f, (size_t)Eo * 4);
    float* out_w = slurp(f, (size_t)Eo * 4);
    int64_t* in_off = slurp(f, (size_t)(V + 1) * 8);
    int32_t* in_nbr = slurp(f, (size_t)Ei * 4);
    float* in_w = slurp(f, (size_t)Ei * 4);
    fclose(f);

    CK(ge_set_graph_csr(ctx, V, out_off, out_nbr, out_w, in_off, in_nbr, in_w));
    int64_t nnz = 0; double xmax = 0;
    CK(ge_cooc_build(ctx, &nnz, &xmax));
    if (nnz != ge_cooc_count(ctx) || V != ge_vocab_size(ctx) || xmax != ge_cooc_max(ctx)) { fprintf(stderr, "accessor mismatch\n"); return 1; }
    int32_t* I = malloc((size_t)nnz * 4); int32_t* J = malloc((size_t)nnz * 4); float* X = malloc((size_t)nnz * 4);
    CK(ge_cooc_export(ctx, I, J, X, nnz));
    double final_cost = -1; int32_t epochs = 0;
    CK(ge_train(ctx, on_epoch, NULL, &final_cost, &epochs));
    if (epochs != epochs_seen) { fprintf(stderr, "callback saw %d epochs, ge_train reports %d\n", epochs_seen, epochs); return 1; }
    double* hist = malloc(sizeof(double) * (size_t)(epochs > 0 ? epochs : 1));
    if (ge_cost_history(ctx, hist, epochs) != epochs) { fprintf(stderr, "history length\n"); return 1; }
    double* emb = malloc(sizeof(double) * (size_t)V * (size_t)cfg.dim);
    CK(ge_extract(ctx, emb));

    f = fopen(argv[2], "wb");
    if (!f) { perror(argv[2]); return 1; }
    fwrite(&nnz, 8, 1, f); fwrite(&xmax, 8, 1, f);
    fwrite(I, 4, (size_t)nnz, f); fwrite(J, 4, (size_t)nnz, f); fwrite(X, 4, (size_t)nnz, f);
    fwrite(&epochs, 4, 1, f); fwrite(&final_cost, 8, 1, f);
    fwrite(hist, 8, (size_t)epochs, f);
    fwrite(emb, 8, (size_t)V * (size_t)cfg.dim, f);
    fclose(f);
    ge_destroy(ctx);
    printf("ok V=%d nnz=%lld epochs=%d\n", V, (long long)nnz, epochs);
    return 0;
}
