/*
 * ytk_gbdt.h -- C ABI of libytkgbdt.so: ytk-learn's data-parallel (histogram) GBDT tree
 * builder on B200.  This is the drop-in boundary: a JNI shim binds these entry points
 * 1:1 behind the reference's ITreeMaker seam (INTEGRATION.md shows the Java side).
 *
 * The reference has no FFI today; each entry point names the Java interface it
 * replaces (paths relative to the reference root,
 * J/ = src/main/java/com/fenbi/ytklearn/).
 *
 * Conventions: plain C, no exceptions; every function returns 0 on success and a
 * negative ytk_status otherwise, with a message in ytk_gbdt_last_error() (per thread).
 * Host pointers are only read/written during the call; nothing is retained.
 * There is no CPU fallback: without a CUDA device every compute entry point fails
 * with YTK_ERR_CUDA.
 */
#ifndef YTK_GBDT_H
#define YTK_GBDT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    YTK_OK = 0,
    YTK_ERR_ARG = -1,        /* bad argument / call order (CheckUtils.check failures in the reference) */
    YTK_ERR_CUDA = -2,       /* CUDA runtime error, or no device */
    YTK_ERR_CAPACITY = -3,   /* caller-provided output too small */
    YTK_ERR_NCCL = -4,       /* NCCL missing or failed (world > 1 only) */
    YTK_ERR_UNSUPPORTED = -5 /* e.g. a feature with more than 65536 bins */
} ytk_status;

typedef struct ytk_gbdt ytk_gbdt;  /* one per (GPU, row shard): the reference's per-thread tree maker */

/* The fields of GBDTOptimizationParams the tree maker reads
 * (J/param/gbdt/GBDTOptimizationParams.java:49-76), already normalised as its
 * constructor does (:148-154; ytk_gbdt_normalize_max_leaf_cnt). */
typedef struct {
    int32_t max_depth;               /* <0: unlimited                                  */
    int32_t max_leaf_cnt;            /* <=0: unlimited                                 */
    int32_t min_split_samples;       /* <=0: off                                       */
    int32_t grow_policy;             /* 0 = "level", 1 = "loss" (TreeGrowPolicy)       */
    double  min_child_hessian_sum;
    double  min_split_loss;
    double  max_abs_leaf_val;        /* <=0: off                                       */
    double  l1;
    double  l2;
    double  histogram_pool_capacity; /* MB; <=0: pool holds max_leaf_cnt histograms    */
    float   learning_rate;
    int32_t class_num;               /* K = numTreeInGroup                             */
} ytk_gbdt_params;

/* What ITreeMaker.make hands back: Tree/TreeNode/TreeNodeStat as SoA, one slot per node
 * in allocation order (J/data/gbdt/Tree.java:58-73, TreeNode.java:33-48,121-149,
 * DataParallelTreeMaker.java:387-393).  Arrays are caller-allocated with `capacity`
 * slots (2*max_leaf_cnt-1 suffices). */
typedef struct {
    int32_t  capacity;
    int32_t  n_nodes;     /* out */
    int32_t *parent;      /* -1 for the root                                           */
    int32_t *left;        /* -1 for leaves                                             */
    int32_t *right;
    int32_t *fid;         /* splitFeatureIndex                                         */
    int32_t *lo, *hi;     /* splitFeaApproSlotInterval                                 */
    float   *node_info;   /* inner: 0.5f*(lo+hi) in slot space; leaf: nodeValue*lr     */
    float   *loss_chg;    /* TreeNodeStat.lossChg (-inf when the node was not searched)*/
    float   *hess_sum;    /* TreeNodeStat.hessSum                                      */
    int64_t *sample_cnt;  /* TreeNodeStat.nodeSampleCnt (global rows)                  */
} ytk_tree_out;

/* TimeStats (J/data/gbdt/TimeStats.java:30-79), milliseconds, per make() call. */
typedef struct {
    double build_hist;        /* buildHist      (includes the two below)               */
    double compute_hist;      /* buildDetail[0]                                        */
    double communicate_hist;  /* buildDetail[1]                                        */
    double init_stats;        /* initStats                                             */
    double find_best_split;   /* findBestSplit - syncBestSplit                         */
    double sync_best_split;   /* syncBestSplit                                         */
    double partition;         /* resetPosition (not timed separately by the reference) */
    double total;             /* totalTime                                             */
    int32_t hist_miss;        /* histMissCnt (DataParallelTreeMaker.java:506)          */
    int32_t hist_built;       /* nodes whose histogram was built from rows             */
    double  hist_cells;       /* sum over those nodes of local rows x sampled features */
    double  hist_rows;        /* sum over those nodes of local rows                    */
    double  hist_rows_indexed;/* ... of which addressed through the row-id list        */
    double  gpu_hist_ms;      /* CUDA-event time of the histogram-build kernels        */
    int32_t kernel_launches;  /* kernels launched by this call                         */
    int32_t hist_launches;    /* ... of which histogram-build launches (one per batch) */
    int32_t peer_collectives; /* 1: per-level collectives ran over NVLink peer memory, 0: NCCL / one GPU */
    int32_t batches;          /* batches of launches = host waits after the root (a level, a node, or a look-ahead set) */
    int32_t device_allocs;    /* cudaMalloc calls made by this call (histogram slots, grown tables): 0 once warmed up */
    double  max_batch_ms;     /* longest single batch, enqueue to host wake-up (a stall shows up here) */
    double  enqueue_ms;       /* device-driven level loop: host time to queue the whole tree ...       */
    double  host_wait_ms;     /* ... and its one wait for the device                                    */
    int32_t stage_times_from_gpu; /* 1: build_hist .. partition above are CUDA-event times of the stages on the
                                     device (they add up to the tree's device time); 0: host wall-clock per stage */
    int32_t graph_replayed;   /* 1: the tree's launches were replayed from a recorded CUDA graph */
    double  device_span_ms;   /* device-driven level loop: first kernel start to last kernel end on the GPU's timer */
    double  hist_events_ms;   /* CUDA-event time around the histogram launches when they were queued directly (0 in a replay):
                                 gpu_hist_ms is then the kernels' own span on the GPU's timer, and the two agree */
    int32_t rendezvous;       /* world > 1: exchanges between the GPUs in this tree (each waits for the slowest rank) */
    int32_t reserved;
} ytk_time_stats;

/* One best split: SplitInfo (J/data/gbdt/SplitInfo.java:36-70) plus the node totals the
 * reference exchanges beside it (DataParallelTreeMaker.java:543-573). */
typedef struct {
    float   loss_chg;     /* -inf when no candidate                                    */
    int32_t fid, lo, hi;  /* -1 when no candidate                                      */
    double  sum_grad, sum_hess;
    float   root_gain, node_value;
} ytk_split;

const char *ytk_gbdt_last_error(void);

/* GBDTOptimizationParams ctor, data-parallel branch (:148-154). Host only. */
int32_t ytk_gbdt_normalize_max_leaf_cnt(int32_t max_depth, int32_t max_leaf_cnt);

/* DataUtils.avgAssign (J/dataflow/DataUtils.java:193-205): out[2*t] = from, out[2*t+1] = to.
 * Host only; the feature-range ownership of worker t among `workers`. */
int ytk_gbdt_avg_assign(int32_t amount, int32_t workers, int32_t *out);

/* SplitInfo.needReplace reduction used by syncBestSplit
 * (J/data/gbdt/SplitInfo.java:99-104, DataParallelTreeMaker.java:640-653): folds `n`
 * per-worker records (rank order) into *out.  Host only. */
int ytk_gbdt_merge_splits(const ytk_split *records, int32_t n, ytk_split *out);

/* Writes the 128-byte NCCL unique id rank 0 must hand to every ytk_gbdt_create of a
 * world > 1 clique (the rendezvous ytk-mp4j's CommMaster performs,
 * bin/local_optimizer.sh:39).  The transport of those bytes is the host's business. */
int ytk_gbdt_nccl_unique_id(void *out128);

/* Replaces `new DataParallelTreeMaker(comm, params, trainData, position, feaApprData)`
 * (J/optimizer/GBDTOptimizer.java:319-320, ctor DataParallelTreeMaker.java:167-227).
 * One instance per GPU = one mp4j rank; `rank`/`world` as comm.getRank()*threadNum+threadId
 * and slaveNum*threadNum.  nccl_id may be NULL when world == 1. */
int ytk_gbdt_create(const ytk_gbdt_params *params, int32_t device, int32_t rank, int32_t world,
                    const void *nccl_id, ytk_gbdt **out);
int ytk_gbdt_destroy(ytk_gbdt *h);

/* FeatureApprData.getFeaSplitValStartIndex() (J/data/gbdt/FeatureApprData.java:216-218):
 * start[F+1] prefix array of bins per feature; n_rows = trainData.sampleNum of this shard.
 * Feature-range ownership is avgAssign(F, world) as in
 * GBDTDataFlow.handleOtherTrainInfo (J/dataflow/GBDTDataFlow.java:252-263). */
int ytk_gbdt_set_bins(ytk_gbdt *h, const int32_t *start, int32_t F, int64_t n_rows);

/* Sparse rows instead of the dense matrix (BASELINE config 4: 10 000 features at ~1 % density; the
 * reference's GBDTCoreData is dense, J/dataflow/GBDTCoreData.java:164-168, and cannot hold that shape --
 * this is the layout a sparse-aware trainer would hand over).  default_slot[f] is the slot of a
 * feature that a row does not store: the bin of the value missing entries are filled with
 * (FillMissingValue, J/feature/gbdt/missing/).  nnz = stored entries of this shard (capacity).  At
 * most 65536 features.  Everything downstream (make_tree, hooks, scores) is the same; results are
 * those of the dense path on the densified rows. */
int ytk_gbdt_set_bins_sparse(ytk_gbdt *h, const int32_t *start, int32_t F, int64_t n_rows,
                             const int32_t *default_slot, int64_t nnz);

/* CSR chunk: rows [row0, row0+nrows), chunks in row order.  Row r holds entries
 * [row_ptr[r], row_ptr[r+1]) of fids/slots (row_ptr has nrows+1 entries and may start at any
 * offset), sorted by feature, slot = the entry's bin (may equal the default slot). */
int ytk_gbdt_upload_csr(ytk_gbdt *h, const int64_t *row_ptr, const int32_t *fids, const int32_t *slots,
                        int64_t row0, int64_t nrows);

/* The same with original values instead of slots (float bits, as GBDTCoreData.x holds them before
 * binning): each stored value is binned on the device with the rule of
 * FeatureApprData.convertFeaVal2ApprFeaIndex (A1, J/data/gbdt/FeatureApprData.java:179-205) against
 * cuts[start[f] .. start[f+1]).  default_slot stays what the caller computed for the fill value. */
int ytk_gbdt_quantize_csr(ytk_gbdt *h, const int64_t *row_ptr, const int32_t *fids, const int32_t *value_bits,
                          int64_t row0, int64_t nrows, const float *cuts);

/* One chunk of GBDTCoreData.x after binning (J/dataflow/GBDTCoreData.java:164-168,
 * FeatureApprData.java:202): int32 slot indices, row-major with stride S = maxFeatureDim.
 * The device keeps them as uint8, row-major, padded to 32 features (uint16 columns for the
 * features with more than 256 bins). */
int ytk_gbdt_upload_rows(ytk_gbdt *h, const int32_t *x_chunk, int64_t row0, int64_t nrows,
                         int32_t stride);
/* The same rows as one byte per slot ("direct u8 upload"): for a trainer that keeps
 * GBDTCoreData.x narrower than int32 once FeatureApprData has binned it (every feature at most 256
 * bins) -- a quarter of the bytes over PCIe.  Same checks, same result as ytk_gbdt_upload_rows. */
int ytk_gbdt_upload_rows_u8(ytk_gbdt *h, const uint8_t *x_chunk, int64_t row0, int64_t nrows, int32_t stride);

/* Same rows, but still holding Float.floatToRawIntBits(value)
 * (GBDTCoreData.java:381): applies FeatureApprData.convertFeaVal2ApprFeaIndex
 * (J/data/gbdt/FeatureApprData.java:179-205, BinarySearch.findLastEqualOrUpper
 * J/utils/BinarySearch.java:33-56) on the device.  cuts = all features' sorted bin
 * centres concatenated, feature f at [start[f], start[f+1]). */
int ytk_gbdt_quantize_rows(ytk_gbdt *h, const int32_t *xbits_chunk, int64_t row0, int64_t nrows,
                           int32_t stride, const float *cuts);

/* Reads back rows [row0, row0+nrows) of the device bin matrix as int32 [nrows][F]. */
int ytk_gbdt_download_rows(ytk_gbdt *h, int32_t *out, int64_t row0, int64_t nrows);

/* One chunk of GBDTCoreData.gradPairs (J/dataflow/GBDTCoreData.java:194-202):
 * float (g,h) at [(row*K + gid)*2 + {0,1}], rewritten by the Java side every round
 * (J/optimizer/GBDTOptimizer.java:550-552,598-599). */
int ytk_gbdt_set_grads(ytk_gbdt *h, const float *gp_chunk, int64_t row0, int64_t nrows);

/* ITreeMaker.make(groupId) (J/optimizer/gbdt/ITreeMaker.java:34-39,
 * DataParallelTreeMaker.make :229-295).  fea_index = the sorted sampled feature ids of
 * initAssistData (:425-469), NULL = all; row_mask (N bytes, 0 = not sampled) replaces the
 * unseeded subsampling of :406-419, NULL = all rows.  ts may be NULL. */
int ytk_gbdt_make_tree(ytk_gbdt *h, int32_t gid, const int32_t *fea_index, int32_t n_fea,
                       const uint8_t *row_mask, ytk_tree_out *out, ytk_time_stats *ts);

/* The same trainer with its `world` row shards on ONE device: `world` handles that share a stream; every
 * collective of the data-parallel scheme (J/data/gbdt/HistogramBuilder.java:95,
 * J/optimizer/gbdt/DataParallelTreeMaker.java:538,560,640-653) becomes a plain kernel over the members'
 * buffers, the ranks' kernels running one after the other.  Same kernels, same [owner][node][segment]
 * staging layout, feature ownership and record merge as one process per GPU -- for hosts with fewer GPUs
 * than shards, and for testing the multi-rank path on one GPU.  Every member gets set_bins / upload_rows /
 * set_grads like any handle; ytk_gbdt_make_tree_clique replaces the `world` concurrent ITreeMaker.make
 * calls (row_masks: NULL or one mask pointer per rank, each NULL or n_rows of that rank; outs/ts: one per
 * rank, ts may be NULL).  Takes what the device-planned level loop takes (level-wise growth, max_depth
 * 1..12, no bounded pool, no row-count stop rule, dense rows); anything else: YTK_ERR_UNSUPPORTED. */
int ytk_gbdt_create_clique(const ytk_gbdt_params *params, int32_t device, int32_t world, ytk_gbdt **out /* [world] */);
int ytk_gbdt_make_tree_clique(ytk_gbdt **members, int32_t world, int32_t gid, const int32_t *fea_index, int32_t n_fea,
                              const uint8_t *const *row_masks, ytk_tree_out *outs, ytk_time_stats *ts);

/* Final partition of the last make(): node id of every local row, -1 if not sampled
 * (SamplePositionData.getPosAndSampleIndex, J/data/gbdt/SamplePositionData.java:64-74;
 * consumed by TreeRefiner, J/optimizer/gbdt/TreeRefiner.java:104-123). */
int ytk_gbdt_get_positions(ytk_gbdt *h, int32_t *node_of_row);

/* ---- per-stage hooks (parity tests and kernel benches; same kernels make() uses) ---- */

/* HistogramBuilder.buildHist (J/data/gbdt/HistogramBuilder.java:56-98) for the listed
 * local rows (NULL = all rows, contiguous): out[2*B] doubles, (g,h) interleaved at
 * 2*(start[fid]+slot); world > 1: reduced over ranks, every rank gets the full array. */
int ytk_gbdt_build_hist(ytk_gbdt *h, int32_t gid, const int32_t *rows, int64_t n_rows,
                        const int32_t *fea_index, int32_t n_fea, double *out);

/* HistogramBuilder.substract (:101-107): out = parent - small over 2*B doubles, computed
 * on the device in the integer domain the tree maker uses. */
int ytk_gbdt_subtract_hist(ytk_gbdt *h, const double *parent, const double *small, double *out);

/* updateTreeMakerNodeStats + enumerateSplit + SplitInfo.update over all sampled features
 * (DataParallelTreeMaker.java:543-637) on a caller-supplied histogram (2*B doubles). */
int ytk_gbdt_find_split(ytk_gbdt *h, const double *hist, int64_t sample_cnt,
                        const int32_t *fea_index, int32_t n_fea, ytk_split *out);

/* SamplePositionData.resetPosition (J/data/gbdt/SamplePositionData.java:115-133) on a
 * caller-supplied row list: stable split into rows with (float)slot(row,fid) < cond
 * (first) and the rest; returns the number of left rows. */
int ytk_gbdt_partition_rows(ytk_gbdt *h, const int32_t *rows, int64_t n_rows, int32_t fid,
                            float cond, int32_t *out_rows, int64_t *left_cnt);

/* ---- the text formats either side of the tree maker (host only, no GPU needed) ------
 * The model text of GBDTModel.dumpModel / loadModel (J/data/gbdt/GBDTModel.java:63-97) and Tree.dumpModel / loadModel
 * (J/data/gbdt/Tree.java:192-291; node patterns :47-48; TreeNodeStat.print J/data/gbdt/TreeNodeStat.java:71-79),
 * byte for byte: floats are printed as java.lang.Float.toString prints them.  A tree is its ytk_tree_out arrays
 * (node_info = leaf values) plus per node: the real-valued threshold of an inner node (convertModel,
 * J/optimizer/GBDTOptimizer.java:663-677), its default child (NULL: the left one) and -- optionally -- feature
 * names indexed by fid (NULL: the decimal fid).  The dump functions return the number of bytes the text needs
 * (without the terminating NUL) and write at most cap bytes; -1 on a malformed tree. */
int ytk_gbdt_float_to_string(float v, char *out, int32_t cap);
int64_t ytk_gbdt_dump_model_header(float base_prediction, int32_t class_num, const char *loss_function, int32_t tree_num,
                                   char *out, int64_t cap);
int64_t ytk_gbdt_dump_tree(const ytk_tree_out *tree, const float *threshold, const int32_t *default_child,
                           const char *const *feature_names, int32_t iter, int32_t with_stats, char *out, int64_t cap);
/* *consumed = bytes of `text` read; ytk_gbdt_load_tree fills the caller's arrays (tree->capacity nodes; fid = the
 * feature name when it is a decimal number, else -1; names: name_stride bytes per node, may be NULL). */
int ytk_gbdt_load_model_header(const char *text, int64_t len, float *base_prediction, int32_t *class_num,
                               char *loss_function, int32_t loss_cap, int32_t *tree_num, int64_t *consumed);
int ytk_gbdt_load_tree(const char *text, int64_t len, ytk_tree_out *tree, float *threshold, int32_t *default_child,
                       char *names, int32_t name_stride, int64_t *consumed);
/* The ytklearn training-row text (docs/data_format.md:1-40: weight x_delim labels x_delim name:value,... [x_delim init
 * prediction]) parsed as GBDTCoreData does (J/dataflow/GBDTCoreData.java:205-249,352-404, J/dataflow/CoreData.java:
 * 369-378) into weight[row], y[row][class_num] (a single class index becomes a one-hot row), init_score (may be NULL) and
 * CSR rows for ytk_gbdt_quantize_csr: row_ptr[rows + 1], fids, value_bits = Float.floatToRawIntBits(value), entries
 * sorted by feature, the last value of a repeated name kept (it is a HashMap there).  dict_sorted: the feature names in
 * lexicographic order = their indices (J/dataflow/DataFlow.java:299-329); names not in it are skipped; NULL: the name
 * is its own decimal index.  Delimiters are literal strings.  Stops at max_rows / max_entries or at an unterminated
 * last line; *consumed = bytes parsed, so a caller streams a file chunk by chunk. */
int ytk_gbdt_parse_rows(const char *text, int64_t len, const char *x_delim, const char *y_delim, const char *features_delim,
                        const char *name_val_delim, int32_t class_num, const char *const *dict_sorted, int32_t n_dict,
                        int64_t max_rows, int64_t max_entries, float *weight, float *y, float *init_score, int64_t *row_ptr,
                        int32_t *fids, int32_t *value_bits, int64_t *n_rows, int64_t *n_entries, int64_t *consumed);

/* ---- TreeRefiner for loss_function = l1 --------------------------------------
 * TreeRefiner.refine (J/optimizer/gbdt/TreeRefiner.java:209-255, exact data-parallel variant :104-123 with
 * PreciseQuantile's weighted median, J/utils/PreciseQuantile.java:156-208): right after make(), every leaf
 * of `tree` (the tree that make() just returned) gets the value learning_rate x weighted median of
 * `residual` over the leaf's rows of ALL ranks -- the sampled rows when a row mask was given.
 *   residual  label - score per local row (TreeRefiner.java:226), or NULL: y - score as
 *             ytk_gbdt_set_targets / ytk_gbdt_update_scores hold them on the device;
 *   weight    per local row, or NULL: GBDTCoreData.weight from set_targets when residual is NULL, else 1.
 * The refined values also replace the leaf values ytk_gbdt_update_scores adds.  world > 1: a collective
 * call (every rank gathers all ranks' pairs over NCCL).  The sketch variant (lad_refine_appr = true,
 * :156-207) is not built: the exact one is what the sketch approximates. */
int ytk_gbdt_refine_leaves(ytk_gbdt *h, const float *residual, const float *weight, ytk_tree_out *tree);

/* ---- the step either side of make(), kept on the device -------------------
 * GBDTOptimizer.predictAndCalcLossGrad, train branch (J/optimizer/GBDTOptimizer.java:513-609): score
 * update from the trees just made, loss, and the next round's gradient pairs, without the (g,h) arrays
 * crossing PCIe every round.  Optional: a trainer that keeps computing gradients in Java simply keeps
 * calling ytk_gbdt_set_grads. */
typedef enum { YTK_LOSS_SIGMOID = 0, YTK_LOSS_L2 = 1, YTK_LOSS_SOFTMAX = 2, YTK_LOSS_L1 = 3 } ytk_loss;

/* GBDTCoreData.y / weight / initScore (J/dataflow/GBDTCoreData.java:286-296,424-447): y has N*K floats
 * (K > 1: one-hot rows), weight N floats (NULL = 1), init_score N*K floats (NULL = 0).  Resets the
 * device score buffer to 0 (a new training run). */
int ytk_gbdt_set_targets(ytk_gbdt *h, int32_t loss, const float *y, const float *weight, const float *init_score);

/* score(data, sid, gid, ...) (J/optimizer/GBDTOptimizer.java:641-658) for the tree the last
 * ytk_gbdt_make_tree(gid) built: score[row*K+gid] += leaf value (float).  Sampled rows take the leaf
 * the partition left them in; rows excluded by row_mask walk the tree on their slots (`<=`,
 * J/data/gbdt/Tree.java:156-168). */
int ytk_gbdt_update_scores(ytk_gbdt *h, int32_t gid);

/* The loss/gradient loop of predictAndCalcLossGrad (:536-607) with SigmoidFunction / L2Function /
 * SoftmaxFunction (J/loss/SigmoidFunction.java, L2Function.java, SoftmaxFunction.java): rewrites the device gradient pairs (as ytk_gbdt_set_grads would)
 * and returns sum_i w_i * loss_i over the local rows. */
int ytk_gbdt_compute_grads(ytk_gbdt *h, double *loss_sum);

/* Read back the device score (N*K floats) / gradient pairs (N*K*2 floats). */
int ytk_gbdt_get_scores(ytk_gbdt *h, float *out);
int ytk_gbdt_get_grads(ytk_gbdt *h, float *out);

/* ---- test-set scoring -----------------------------------------------------
 * GBDTOptimizer.score with isOriginTree = true (J/optimizer/GBDTOptimizer.java:630-650) through
 * Tree.getLeafIndex / getNext (J/data/gbdt/Tree.java:142-168): rows hold ORIGINAL feature values as
 * float bits (GBDTCoreData.x before binning, int32 row-major with `stride` per row); value <=
 * threshold goes left; a value equal to Constants.INT_MISSING_VALUE (the raw bits of Float.NaN,
 * 0x7fc00000) takes default_child[node] (null: the left child).  `tree` is what make_tree returned;
 * threshold[node] is the real-valued split of inner nodes after
 * Tree.convertFeatureSplitValueInModel (:293-309).  score[row*class_num + gid] += leaf value (float),
 * in place on the host array.  Stateless: any rows, not the training shard. */
int ytk_gbdt_score_values(ytk_gbdt *h, const int32_t *xbits, int64_t nrows, int32_t stride,
                          const ytk_tree_out *tree, const float *threshold, const int32_t *default_child,
                          int32_t class_num, int32_t gid, float *score);

/* No counterpart in the reference: an addition for hosts that must not allocate device memory
 * (all of it belongs to their own pool, or the hidden device-wide synchronisation of
 * cudaMalloc/cudaFree is not acceptable where they call from).  Same operation and same kernels
 * as ytk_gbdt_partition_rows, but stateless and on DEVICE buffers the caller already owns:
 *   d_bins      uint8 slots, row-major, `stride` bytes per row
 *   d_rows_in   n_rows row ids;  d_rows_out  n_rows ids, left rows first, both sides in input order
 *   d_left_cnt  one int64: number of rows with (float)slot(row,fid) < cond
 * Work is queued on `cuda_stream` (a cudaStream_t, NULL = default stream); results are valid once
 * that stream has reached this point.  The call itself calls no cudaMalloc/cudaFree: its work space
 * is a fixed array inside the library, 4 MiB + 17 KiB held on every GPU that loads the library
 * until the process exits.  Hence:
 *   - at most 4,194,304 rows per call, more is refused with YTK_ERR_CAPACITY;
 *   - the array is shared by all calls in the process: two calls that can run on the same GPU at
 *     the same time (different streams or threads) race on it -- keep them ordered;
 *   - the first call on each GPU also looks up the array's device address (the device that is
 *     current at the call); make that first call before any stream capture;
 *   - `cuda_stream` must belong to the device that is current at the call (the array used is that
 *     device's), and d_rows_in / d_rows_out must not overlap (refused with YTK_ERR_ARG). */
int ytk_gbdt_partition_rows_static(const uint8_t *d_bins, int32_t stride, const int32_t *d_rows_in,
                                   int64_t n_rows, int32_t fid, float cond, int32_t *d_rows_out,
                                   int64_t *d_left_cnt, void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif
