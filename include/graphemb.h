/* graphemb.h — C ABI of libgraphemb.so, the B200-native hot path of Phaken/graph-embeddings.
 *
 * The reference has no FFI layer; its seam is two Java interfaces consumed by
 * Main.runProgram (J/Main.java:70-77, J = src/main/java/org/uu/nl/embedding/):
 *     CoOccurrenceMatrix  (J/util/CoOccurrenceMatrix.java:6-16)  <- new BookmarkColoring(graph, config)
 *     IOptimizer/Optimum  (J/opt/IOptimizer.java:6-11, J/opt/Optimum.java:9-41) <- optimizer.optimize()
 * Each entry point below names the reference code it replaces.  INTEGRATION.md shows the
 * Java-side binding (JNI for the Java 8 target of pom.xml:13-14, Panama FFM on JDK >= 22).
 *
 * Conventions: plain C, opaque handle, caller-owned buffers, int status (0 = OK, negative =
 * error, text via ge_last_error).  Never throws, never aborts, never keeps a caller pointer
 * after return.  One call at a time per handle.  Callbacks run on the calling thread.
 * All host pointers are ordinary host memory unless a function says "device".
 */
#ifndef GRAPHEMB_H
#define GRAPHEMB_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GE_OK 0
#define GE_ERR_INVALID (-1)   /* bad argument / configuration (Configuration.check, :478-494) */
#define GE_ERR_STATE (-2)     /* call order violated (e.g. train before a matrix exists)       */
#define GE_ERR_CUDA (-3)      /* CUDA runtime error; message has the cudaError string          */
#define GE_ERR_NOMEM (-4)     /* device or host allocation failed                              */
#define GE_ERR_GRAPH (-5)     /* neighbour lists not sorted/unique, ids out of range           */
#define GE_ERR_DIST (-6)      /* NCCL / multi-GPU set-up failure                               */
#define GE_ERR_NODEVICE (-7)  /* no CUDA device: there is NO CPU fallback                      */

/* Configuration.EmbeddingMethod (J/util/config/Configuration.java:19-21) */
enum { GE_METHOD_GLOVE = 0, GE_METHOD_PGLOVE = 1 };
/* Configuration.OptimizationMethod (:23-25): J/opt/grad/Adagrad.java, Adam.java, AMSGrad.java.
 * The three shipped *.config.yml select adagrad; adam / amsgrad are SURVEY §8(f) row 1. */
enum { GE_OPT_ADAGRAD = 0, GE_OPT_ADAM = 1, GE_OPT_AMSGRAD = 2 };
/* Configuration.BCANormalization (:31-33) */
enum { GE_NORM_NONE = 0, GE_NORM_UNITY = 1, GE_NORM_COUNTS = 2 };
/* Arithmetic of the AdaGrad step.
 * GE_MATH_FAST : f32 arithmetic (rsqrt), one warp per nonzero, Hogwild across warps — the
 *                throughput path.
 * GE_MATH_JAVA : the reference's exact promotion pattern (fp64 divide/sqrt/log/pow islands,
 *                J/opt/grad/Adagrad.java:58-93).  With ge_config.serial = 1 it runs one warp
 *                in the reference's single-thread order (the parity trajectory). */
enum { GE_MATH_FAST = 0, GE_MATH_JAVA = 1 };
/* Epoch order.  GE_ORDER_JAVA replays java.util.Random's cumulative Fisher-Yates on the host
 * (J/util/rnd/Permutation.java:21-23, ExtendedRandom.java:398-407); GE_ORDER_DEVICE uses a
 * keyed bijection evaluated on the device (no permutation array). */
enum { GE_ORDER_DEVICE = 0, GE_ORDER_JAVA = 1 };
/* Store initialisation.  GE_INIT_JAVA replays the LCG stream of Optimizer.java:50-57 bit for
 * bit (on the device, by LCG jump-ahead); GE_INIT_COUNTER is a counter-based generator with
 * the same distribution (U[0,1)-0.5)/dim. */
enum { GE_INIT_COUNTER = 0, GE_INIT_JAVA = 1 };

/* How the FAST path writes a row back.  GE_UPDATE_ATOMIC: deltas are applied with vector
 * red.global.add.f32 at the L2 — concurrent warps on one hub row never lose each other's
 * update (only the gradient is stale, as in the reference's own Hogwild threads).
 * GE_UPDATE_STORE: plain read-modify-write stores, exactly the reference's racy pattern
 * (Adagrad.java:76-80) but with thousands of warps instead of nproc-1 threads. */
enum { GE_UPDATE_ATOMIC = 0, GE_UPDATE_STORE = 1, GE_UPDATE_ATOMIC_ADAM = 2 };
/* adam / amsgrad always use plain stores unless GE_UPDATE_ATOMIC_ADAM (experiment: red.add of the
 * parameter deltas; overshoots on hub rows, see csrc/train.cu) is selected. */

typedef struct ge_config {
    int32_t dim;           /* Configuration.dim                                   */
    int32_t method;        /* GE_METHOD_*   <- Configuration.method               */
    int32_t opt_method;    /* GE_OPT_*      <- Configuration.opt.method           */
    int32_t directed;      /* bca.directed (reverse pass is implied when directed) */
    int32_t normalize;     /* GE_NORM_*     <- bca.normalize                      */
    double alpha;          /* bca.alpha                                           */
    double epsilon;        /* bca.epsilon                                         */
    double tolerance;      /* opt.tolerance                                       */
    int32_t maxiter;       /* opt.maxiter                                         */
    int64_t seed;          /* reference: System.currentTimeMillis() (Main.java:62) */
    float learning_rate;   /* hard-wired 0.05f in Optimizer.java:26; 0 -> 0.05f   */
    int32_t math;          /* GE_MATH_*                                           */
    int32_t order;         /* GE_ORDER_*                                          */
    int32_t init;          /* GE_INIT_*                                           */
    int32_t serial;        /* 1: single warp, reference single-thread order       */
    int32_t device;        /* CUDA device ordinal                                 */
    int32_t update;        /* GE_UPDATE_*                                         */
} ge_config;

typedef struct ge_ctx ge_ctx;

/* per-epoch progress (Optimizer.java:97-101): return non-zero to stop early */
typedef int (*ge_epoch_cb)(void* user, int32_t epoch, double cost, double seconds);

/* fills *cfg with the values of the shipped *.config.yml files: pglove, dim 300, alpha .1,
 * epsilon 1e-3, directed, adagrad, tolerance 1e-4, maxiter 1000 (dblp.config.yml:2-3,17-26) */
void ge_config_default(ge_config* cfg);

int ge_create(const ge_config* cfg, ge_ctx** out);
void ge_destroy(ge_ctx* ctx);
const char* ge_last_error(const ge_ctx* ctx); /* ctx may be NULL: last create error */
const char* ge_version(void);
/* run all device work of this handle on the caller's CUDA stream (a cudaStream_t passed as
 * void*; NULL restores the handle's own stream) so a host that owns streams/events (e.g. a
 * torch process timing with CUDA events) brackets exactly this library's launches */
int ge_set_stream(ge_ctx* ctx, void* cuda_stream);

/* A1 — adjacency hand-off (BookmarkColoring.java:49-52).  Neighbour lists must be sorted by
 * id and unique; off arrays have V+1 entries.  Copies to the device. */
int ge_set_graph_csr(ge_ctx*, int32_t V, const int64_t* out_off, const int32_t* out_nbr,
                     const float* out_w, const int64_t* in_off, const int32_t* in_nbr,
                     const float* in_w);

/* SURVEY §8(f) row 3 — the same hand-off computed ON THE DEVICE from the raw directed edge list
 * (replaces getIn/OutNeighborhoods + In/OutEdgeNeighborhoodAlgorithm.compute,
 * J/convert/util/EdgeNeighborhoodAlgorithm.java:42-77).  Parallel edges collapse to one neighbour
 * whose weight is that of the first such edge in input order; self-loops are kept. */
int ge_set_graph_edges(ge_ctx*, int32_t V, int64_t E, const int32_t* src, const int32_t* dst, const float* w);
/* sizes of the CSR now on the device (unique out- / in-neighbour entries), seconds spent flattening */
int ge_graph_sizes(ge_ctx*, int64_t* E_out, int64_t* E_in, double* flatten_seconds);
/* copy the device CSR back (validation); any pointer may be NULL */
int ge_graph_export(ge_ctx*, int64_t* out_off, int32_t* out_nbr, float* out_w, int64_t* in_off,
                    int32_t* in_nbr, float* in_w);

/* SURVEY §8(f) row 4 — similarity edges between literals: the all-pairs compare that
 * Rdf2GrphConverter.convert runs per predicate group just before the graph is handed over
 * (J/convert/Rdf2GrphConverter.java:127-186, one CompareJob per source literal,
 * J/compare/CompareJob.java:32-53: every target j >= startIndex whose vertex differs, pair kept when
 * metric.similarity(s1, s2) >= threshold, weight (float) similarity, later added as TWO directed
 * edges of type 0).  Strings stay on the host: the caller hands over what the metrics read —
 * profile-based metrics (J/util/similarity/PreComputed*.java) get each literal's profile as
 * (shingle id, count) lists sorted by id, NUMERIC / DATE_* get the parsed values,
 * JAROWINKLER / LEVENSHTEIN (info.debatty java-string-similarity, Configuration.java:212-215) get the
 * labels themselves as UTF-16 code units. */
#define GE_SIM_NGRAM_COSINE 0  /* PreComputedNgramCosine.java:20-31,55-57  */
#define GE_SIM_NGRAM_JACCARD 1 /* PreComputedNgramJaccard.java:28-47       */
#define GE_SIM_TOKEN_COSINE 2  /* PreComputedTokenCosine.java:29-44        */
#define GE_SIM_TOKEN_JACCARD 3 /* PreComputedTokenJaccard.java:32-51       */
#define GE_SIM_NUMERIC 4       /* Numeric.java:18-44                       */
#define GE_SIM_DATE_DAYS 5     /* Date.java:30-66, DateDays/Months/Years   */
#define GE_SIM_DATE_MONTHS 6
#define GE_SIM_DATE_YEARS 7
#define GE_SIM_JAROWINKLER 8   /* info.debatty JaroWinkler.similarity (threshold 0.7, float Jaro term) */
#define GE_SIM_LEVENSHTEIN 9   /* info.debatty NormalizedLevenshtein.similarity = 1 - distance / max length */
#define GE_SIM_TIME_BIDIRECTIONAL 0 /* Configuration.SimilarityGroup.Time (Configuration.java:184-186) */
#define GE_SIM_TIME_BACKWARDS 1
#define GE_SIM_TIME_FORWARDS 2
#define GE_SIM_FLAG_EMPTY 1  /* label.isEmpty()                                        */
#define GE_SIM_FLAG_LONG 2   /* label.length() >= ngram (NGRAM_COSINE, :27)            */
#define GE_SIM_FLAG_PARSED 4 /* DATE_*: LocalDate.parse succeeded (else similarity 0)  */
#define GE_SIM_NUM_OK 0
#define GE_SIM_NUM_NFE 1 /* Integer.parseInt threw NumberFormatException -> similarity 0 (Numeric.java:40-43)    */
#define GE_SIM_NUM_OOB 2 /* substring(0, cut) past the end: uncaught, the whole CompareJob of s1 is lost (:33) */

typedef struct ge_sim_group {
    int32_t method;         /* GE_SIM_* */
    int32_t upper_triangle; /* sourcePredicate.equals(targetPredicate) (Rdf2GrphConverter.java:50): j starts at index+1 */
    double threshold;       /* pair kept when similarity >= threshold (CompareJob.java:46) */
    double smooth;          /* getSmooth(): Numeric alpha / Date smooth, 0 already mapped to 1 */
    double distance;
    int32_t time_dir;       /* GE_SIM_TIME_* (DATE_* only) */
    int32_t n_items;        /* distinct literal vertices of the group (source and target sets together) */
    const int32_t* vert;      /* [n_items] graph vertex id                                      */
    const int32_t* string_id; /* [n_items] literals with equal labels share an id (s1.equals(s2)) */
    const uint8_t* flags;     /* [n_items] GE_SIM_FLAG_*                                        */
    /* profile metrics: item k's profile is prof_key/prof_cnt[prof_off[k] .. prof_off[k+1]), keys ascending */
    const int64_t* prof_off;
    const int32_t* prof_key;
    const int32_t* prof_cnt;
    /* NUMERIC.  Numeric.java:29-33 cuts BOTH strings at s1's '^' position, so what s2 parses to
     * depends on s1: the caller lists the n_cut distinct cut positions of the group (one of them may
     * stand for "no '^': parse the whole label"), cut_slot[k] is the slot of item k's own position,
     * and num_val / num_status[k * n_cut + c] are Integer.parseInt of item k's label cut at slot c. */
    int32_t n_cut;
    const int32_t* cut_slot;
    const int32_t* num_val;
    const uint8_t* num_status; /* GE_SIM_NUM_* */
    /* DATE_*: LocalDate fields of the parsed labels */
    const int64_t* epoch_day;    /* LocalDate.toEpochDay()                       */
    const int64_t* packed_month; /* getProlepticMonth() * 32 + getDayOfMonth()   */
    /* the two sides, as item indices in the iteration order of group.source / group.target */
    int32_t n_source;
    const int32_t* source;
    int32_t n_target;
    const int32_t* target;
    /* JAROWINKLER / LEVENSHTEIN: item k's label is chars[char_off[k] .. char_off[k+1]) (UTF-16 code units) */
    const int64_t* char_off;
    const uint16_t* chars;
} ge_sim_group;

/* runs the compare on the device and keeps the result there; count = pairs kept.  Pairs are
 * ordered by (source index, target index) — the reference's order depends on which job finishes
 * first. */
int ge_similarity_pairs(ge_ctx*, const ge_sim_group* group, int64_t* count, double* seconds);
/* copy the kept pairs out: vert[k] = source vertex, other[k] = target vertex, sim[k] = (float) similarity */
int ge_similarity_export(ge_ctx*, int64_t cap, int32_t* vert, int32_t* other, float* sim);

/* A2-A4 — new BookmarkColoring(graph, config): builds the COO on the device.
 * Outputs: nnz = coOccurrenceCount(), xmax = max().  Rows [first, first+count) only when
 * count >= 0 (multi-GPU row sharding); count < 0 = all rows. */
int ge_cooc_build(ge_ctx*, int64_t* nnz, double* xmax);
int ge_cooc_build_rows(ge_ctx*, int32_t first, int32_t count, int64_t* nnz, double* xmax);
/* counters of the last build: pops (pair events), scanned adjacency entries, seconds on device */
int ge_cooc_stats(ge_ctx*, int64_t* pops, int64_t* scanned, double* seconds);
/* adjacency entries the last build really read: `scanned` counts every entry the reference's loop visits
 * (DirectedWeighted.java:79-98); lists whose heaviest edge cannot reach epsilon are skipped unread here.
 * tier2_rows: bookmarks whose record table outgrew shared memory and were re-run on a global-memory table */
int ge_cooc_fetched(ge_ctx*, int64_t* fetched, int64_t* tier2_rows);

/* A6 — cIdx_I/J/C for all k (unpermuted); cap = capacity of the three arrays. */
int ge_cooc_export(ge_ctx*, int32_t* I, int32_t* J, float* X, int64_t cap);
/* feed an external COO (validation; C5-style direct COO).  xmax <= 0: computed. */
int ge_cooc_import(ge_ctx*, int32_t V, int64_t nnz, const int32_t* I, const int32_t* J,
                   const float* X, double xmax);
int64_t ge_cooc_count(ge_ctx*);   /* coOccurrenceCount() */
int32_t ge_vocab_size(ge_ctx*);   /* vocabSize() */
double ge_cooc_max(ge_ctx*);      /* max() */

/* B1 — allocate + initialise the embedding store (Optimizer ctor :45-57, Adagrad ctor :22-33).
 * Called implicitly by ge_train when missing. */
int ge_store_init(ge_ctx*);
/* B4/B5 — one epoch: shuffle + all nonzeros; cost = localCost (Optimizer.java:79-96). */
int ge_train_epoch(ge_ctx*, double* cost);
/* B6 — optimizer.optimize(): epoch loop with the tolerance stop rule (:67-120). */
int ge_train(ge_ctx*, ge_epoch_cb cb, void* user, double* final_cost, int32_t* epochs);
int ge_cost_history(ge_ctx*, double* out, int32_t cap); /* Optimum.costHistory; returns count */
/* B7 — extractResult(): (focus+context)/2, row-major V x dim. */
int ge_extract(ge_ctx*, double* out);
int ge_extract_f32(ge_ctx*, float* out);
/* raw store rows in the reference's layout (validation): which = 0 focus, 1 context,
 * 2 fBias, 3 cBias, 4 gradSqFocus, 5 gradSqContext, 6 gradSqFBias, 7 gradSqCBias (adam / amsgrad:
 * the second moments M2*), 8 M1focus, 9 M1context, 10 M1fBias, 11 M1cBias (adam / amsgrad only) */
int ge_store_export(ge_ctx*, int32_t which, float* out);
/* epoch order used by the last/next epoch as an explicit permutation (validation) */
int ge_epoch_order(ge_ctx*, int32_t epoch, int32_t* perm, int64_t cap);
/* forget the packed nonzeros and the embedding store (the matrix stays): the next epoch re-packs and
 * starts from a fresh initialisation — a new Optimizer on the same CoOccurrenceMatrix (Main.java:72) */
int ge_train_reset(ge_ctx*);
/* device seconds and kernel launches spent in update kernels so far */
int ge_train_stats(ge_ctx*, double* update_seconds, int64_t* launches, int64_t* updates);
/* Hub report of the prepared trainer (single-GPU or block schedule).  The reference has at most
 * nproc-1 writers on an embedding row (Optimizer.java:84-92); a GPU wave has 4 736 warps.  Ids that
 * would see more than GE_HOT_WRITERS (16) of them get damped deltas; when the damped ids hold GE_HUB_MASS
 * (5 %) or more of a launch's nonzeros, the launch is capped to the warp count that puts GE_HUB_SERIAL
 * (8; half of it in the buckets of the block schedule) on the hottest id.
 *   max_concurrency : warps of an uncapped wave expected on the hottest id at once
 *   damped_share    : largest share of a launch's nonzeros that touch a damped id
 *   capped_launches : launches (1, or buckets of the block schedule) that run capped
 *   warp_cap        : the (smallest) cap, 0 = none */
int ge_train_hub_info(ge_ctx*, double* max_concurrency, double* damped_share, int32_t* capped_launches, int64_t* warp_cap);

/* SURVEY §8(f) row 2 — EmbeddingTextWriter.write (J/util/write/EmbeddingTextWriter.java:73-149)
 * served from the device store: <vectors_path> gets one line of "%11.6E" values (Java Formatter
 * semantics, see ge_format_e) joined by tabs per emitted vertex, <dict_path> gets "key\tTYPE".
 *   header        : the '#' settings block (:40-70), already formatted by the host; may be NULL
 *   keys          : V C-strings = getKey(i); NULL (or a NULL entry) -> the decimal index
 *   types         : V node types (0 URI, 1 BLANK, 2 LITERAL = NodeInfo ids); NULL -> all URI
 *   type_enabled  : 3 flags = output.outputUri/Blank/LiteralNodes (:36-38); NULL -> all enabled
 *   prefixes      : concatenated prefix lists URI | BLANK | LITERAL, nprefix[3] their lengths;
 *                   an empty list disables the filter for that type (:114-124)
 *   written       : number of vertices emitted */
int ge_write_embeddings(ge_ctx*, const char* vectors_path, const char* dict_path, const char* header,
                        const char* const* keys, const uint8_t* types, const uint8_t* type_enabled,
                        const char* const* prefixes, const int32_t* nprefix, int64_t* written);
/* String.format("%11.6E", v) exactly as java.util.Formatter prints a double (shortest round-trip
 * digits rounded half-up — not C's exact-value rounding).  Returns the length, or -(needed) when
 * cap is too small.  Pure host code: callable without a device. */
int ge_format_e(double v, char* out, int32_t cap);

/* --- multi-GPU (one process per GPU).  No reference counterpart: replaces the
 * thread slicing of J/opt/Optimizer.java:59-63 across devices. ---------------------------- */
#define GE_NCCL_ID_BYTES 128
/* rank 0 creates the id; the host distributes it (torch.distributed / file / env) */
int ge_dist_unique_id(void* id128);
/* world == 1 needs no id: the block schedule then runs on one GPU (rotation = pointer shift).
 * world > 1: the ranks of one node meet in a shared-memory segment named after the id (handles, scalars, barriers)
 * and move data through CUDA IPC peer memory; an NCCL communicator is created only for GE_DIST_ROT=nccl|pipe,
 * GE_DIST_PLUMBING=nccl, or when shared memory / IPC are unavailable. */
int ge_dist_init(ge_ctx*, int32_t rank, int32_t world, const void* id128);
/* this rank's id range in an nblocks x nblocks grid over V ids: build exactly these rows
 * (ge_cooc_build_rows) or import nonzeros whose rows lie in it */
int ge_dist_row_range(ge_ctx*, int32_t V, int32_t nblocks, int32_t* first, int32_t* count);
/* Id layout of the block grid, to be chosen before ge_dist_row_range / the build / ge_dist_prepare.
 * 0 = contiguous (block b = ids [b*Bs, (b+1)*Bs), Bs = ceil(V / nblocks)); 1 = cyclic (block b = ids b,
 * b + nblocks, ...), which balances the buckets of id-ordered graphs (node types in consecutive id ranges, as
 * Rdf2GrphConverter numbers them in encounter order: J/convert/Rdf2GrphConverter.java).  Cyclic: ge_dist_row_range returns a range of VIRTUAL ids
 * vid(v) = (v mod nblocks) * Bs + v div nblocks (whole blocks; a vid whose vertex is >= V is padding); build
 * with ge_cooc_build_rows_cyclic, or import the nonzeros of the rows ge_dist_local_ids lists. */
int ge_dist_set_layout(ge_ctx*, int32_t layout);
int ge_cooc_build_rows_cyclic(ge_ctx*, int32_t nblocks, int32_t first_vid, int32_t count, int64_t* nnz, double* xmax);
/* after ge_dist_prepare: the vertex of every row ge_dist_extract emits, in that order; returns their number */
int64_t ge_dist_local_ids(ge_ctx*, int32_t* ids, int64_t cap);
/* all-reduce max()/nnz, bucket the local nonzeros by (row-block, col-block), allocate and
 * initialise this rank's W rows and its ring of C-block slots */
int ge_dist_prepare(ge_ctx*, int32_t nblocks);
/* one epoch = nblocks sub-epochs, each followed by the ring rotation of one C block
 * (ncclSend to rank-1 / ncclRecv from rank+1); cost = global localCost (all-reduced) */
int ge_dist_train_epoch(ge_ctx*, double* cost);
/* (W+C)/2 of this rank's id range, row-major count x dim (f64: what extractResult() returns, Optimizer.java:130-140) */
int ge_dist_extract_f32(ge_ctx*, float* out, int64_t cap);
int ge_dist_extract(ge_ctx*, double* out, int64_t cap);
/* A1 hand-off for N ranks: one host -> device copy on `root` + ncclBroadcast of the CSR arrays over NVLink instead
 * of N uploads across PCIe.  Non-root ranks pass V = 0 and NULL arrays.  world == 1: same as ge_set_graph_csr. */
int ge_dist_set_graph_csr(ge_ctx*, int32_t root, int32_t V, const int64_t* out_off, const int32_t* out_nbr,
                          const float* out_w, const int64_t* in_off, const int32_t* in_nbr, const float* in_w);
/* data plane of the rotation in use: 2 = half blocks pushed by the copy engine into the neighbour's ring
 * (CUDA IPC mapping, flag hand-shake with stream memory operations; default), 1 = half-block
 * ncclSend/ncclRecv on a second stream (GE_DIST_ROT=pipe, or the fallback when IPC is unavailable),
 * 0 = whole-block ncclSend/ncclRecv in the compute stream (GE_DIST_ROT=nccl), -1 = single GPU */
int ge_dist_rotation_mode(ge_ctx*);
/* returns the number of local buckets and fills their nnz; comm_seconds = device time of this rank's
 * outgoing block copies (copy-engine plane; 0 for the NCCL planes, whose send/recv time includes
 * waiting for the peer) */
int ge_dist_stats(ge_ctx*, int64_t* bucket_nnz, int32_t cap, double* comm_seconds, int64_t* nnz_global);

/* --- allocation-free variant.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
 * that forbid it). ------------------------------------------------------------------------------- */
#define GE_STATIC_MAX_EDGES 1048576        /* largest E the static scratch holds (2^20 edges)          */
#define GE_STATIC_SCRATCH_BYTES 29623808   /* device memory every GPU that loads the library holds     */
/* One direction of ge_set_graph_edges (by_dst = 0: out-lists, major id = src; 1: in-lists, major id = dst)
 * computed by the same kernels, with the sort buffers in a file-scope device array instead of
 * cudaMalloc'ed scratch.  Calls no cudaMalloc / cudaFree / allocating routine and does not synchronise:
 * the work is queued on the handle's stream (ge_set_stream).  ALL data pointers are DEVICE memory
 * of the handle's device, owned by the caller:
 *   d_src, d_dst, d_w : E edges (vertex ids in [0, V), f32 weights)
 *   d_off  [V + 1]    : list offsets
 *   d_nbr, d_wout [E] : neighbours / weights, the first d_info[0] entries are written
 *   d_info [2]        : {unique entries, 1 when a vertex id was out of range (outputs then meaningless)}
 * E > GE_STATIC_MAX_EDGES -> GE_ERR_INVALID (use ge_set_graph_edges).  E == 0 writes zero offsets and is the
 * warm-up call (resolves the scratch address for the device; call it once before a stream capture).
 * Every call on a device shares the one scratch array: calls that overlap on the GPU (other streams,
 * other handles, other threads) race — serialise them. */
int ge_flatten_edges_static(ge_ctx*, int32_t V, int64_t E, int32_t by_dst, const int32_t* d_src,
                            const int32_t* d_dst, const float* d_w, int64_t* d_off, int32_t* d_nbr,
                            float* d_wout, int64_t* d_info);

#ifdef __cplusplus
}
#endif
#endif /* GRAPHEMB_H */
