/*
 * ytk_oracle.h -- CPU restatement of ytk-learn's data-parallel GBDT tree builder.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may build, load or call it, and only as the checker (or as the timed CPU
 * baseline), never as a fallback for the CUDA path.
 *
 * Every function restates one piece of the reference (paths relative to
 * /root/reference, J/ = src/main/java/com/fenbi/ytklearn/).  The reference has no
 * tests and cannot run here (no JVM); parity of this oracle is pinned by the two
 * known-answer outputs the reference prints in its demo READMEs
 * (demo/gbdt/binary_classification/README.md:43-49 and
 * demo/gbdt/multiclass_classification/README.md:53) -- see tests/test_oracle_kat.py.
 */
#ifndef YTK_ORACLE_H
#define YTK_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Fields of GBDTOptimizationParams the tree maker reads
 * (J/param/gbdt/GBDTOptimizationParams.java:49-76), already normalised the way
 * its constructor does (:148-154, see orc_normalize_max_leaf_cnt). */
typedef struct {
    int32_t max_depth;              /* <0: no limit                         */
    int32_t max_leaf_cnt;           /* <=0: no limit                        */
    int32_t min_split_samples;      /* <=0: off                             */
    int32_t grow_policy;            /* 0 = level-wise, 1 = loss-wise        */
    double  min_child_hessian_sum;
    double  min_split_loss;
    double  max_abs_leaf_val;       /* <=0: off                             */
    double  l1;
    double  l2;
    double  histogram_pool_capacity;/* MB, <=0: capacity = max_leaf_cnt     */
    float   learning_rate;
    int32_t class_num;              /* K = numTreeInGroup                   */
} orc_params;

/* SoA tree, one slot per node in allocation order (J/data/gbdt/Tree.java:58-73,
 * TreeNode.java:33-48, DataParallelTreeMaker.java:387-393). */
typedef struct {
    int32_t  capacity;   /* in: slots available in every array below */
    int32_t  n_nodes;    /* out */
    int32_t *parent, *left, *right, *fid, *lo, *hi;
    float   *node_info;  /* inner: 0.5f*(lo+hi) (slot space); leaf: value*lr */
    float   *loss_chg;   /* best.lossChg (-inf when never searched)          */
    float   *hess_sum;   /* (float) sumHess                                  */
    int64_t *sample_cnt;
} orc_tree;

/* J/utils/BinarySearch.java:33-56 + J/data/gbdt/FeatureApprData.java:186-199 */
int32_t orc_find_bin(const float *cuts, int32_t m, float v);

/* FeatureApprData.convertFeaVal2ApprFeaIndex (J/data/gbdt/FeatureApprData.java:179-205):
 * in place over an int32 row-major matrix holding float bit patterns. */
void orc_convert_to_bins(int32_t *x, int64_t n, int32_t S, int32_t F,
                         const float *cuts, const int32_t *start);

/* DataUtils.avgAssign (J/dataflow/DataUtils.java:193-205): out[2*t],out[2*t+1] */
void orc_avg_assign(int32_t amount, int32_t bins, int32_t *out);

/* GBDTOptimizationParams ctor, data-parallel branch (:148-154). */
int32_t orc_normalize_max_leaf_cnt(int32_t max_depth, int32_t max_leaf_cnt);

/* UpdateStrategy.calcGain / calcNodeValue (J/optimizer/gbdt/UpdateStrategy.java:55-100) */
double orc_calc_gain(const orc_params *p, double G, double H);
double orc_calc_node_value(const orc_params *p, double G, double H);

/* HistogramBuilder.buildHist's accumulation loop (J/data/gbdt/HistogramBuilder.java:69-90)
 * for one worker: pos_idx = (nid,row) pairs, interval [v1,v2). hist has 2*B doubles
 * and is zeroed first (:69-71). */
void orc_build_hist(const int32_t *x, int32_t S, const float *gp, int32_t K, int32_t gid,
                    const int32_t *pos_idx, int64_t v1, int64_t v2,
                    const int32_t *fea_index, int32_t n_fea, const int32_t *start,
                    int32_t B, double *hist);

/* DataParallelTreeMaker.enumerateSplit (J/optimizer/gbdt/DataParallelTreeMaker.java:598-637)
 * for one feature on a global histogram (2*B doubles); updates best[] with the
 * SplitInfo.needReplace rule (J/data/gbdt/SplitInfo.java:99-140).
 * best = {lossChg, fid, lo, hi} */
typedef struct { float loss_chg; int32_t fid, lo, hi; } orc_split;
void orc_split_clear(orc_split *s);
int  orc_split_need_replace(const orc_split *cur, float new_loss_chg, int32_t new_fid);
void orc_enumerate_split(const orc_params *p, const double *hist, const int32_t *start,
                         int32_t fid, double sumG, double sumH, float root_gain,
                         orc_split *best);

/* Shapes whose dense int32 matrix does not fit in memory (10 000 features at 1 % density): hold the rows
 * as CSR (entries sorted by feature inside a row) + one default slot per feature and call
 * orc_make_tree with x = NULL; x[row][fid] is then looked up on the fly by the same dense loops.
 * Pass NULLs to clear.  Not thread-safe against a concurrent orc_make_tree. */
void orc_set_virtual_csr(const int64_t *row_ptr, const int32_t *fids, const int32_t *slots, const int32_t *default_slot);

/* SamplePositionData.resetPosition + sortByNodeIndex
 * (J/data/gbdt/SamplePositionData.java:115-165): returns leftCnt. */
int64_t orc_reset_position(int32_t *pos_idx, int64_t v1, int64_t v2, const int32_t *x,
                           int32_t S, int32_t fid, float cond, int32_t left, int32_t right);

/*
 * DataParallelTreeMaker.make (J/optimizer/gbdt/DataParallelTreeMaker.java:229-295) with
 * W emulated workers (mp4j ranks).  Worker w owns rows [row_off[w], row_off[w+1]) of
 * the flat matrix and features avgAssign(F, W)[w]; its histogram is reduced with the
 * others by plain fp64 addition in rank order (stand-in for mp4j reduceScatterArray,
 * HistogramBuilder.java:95).
 *
 *  x        int32 slots, row-major, stride S            (GBDTCoreData.x, flattened)
 *  gp       float (g,h) pairs, [(row*K + gid)*2 + {0,1}] (GBDTCoreData.gradPairs)
 *  row_mask nullable, N bytes: 0 = row not sampled (pos = -1, :406-419)
 *  node_of_row nullable, N ints out: final node id of every row (-1 if unsampled)
 *  hist_miss nullable out: histMissCnt (:506)
 *  threads  >1: worker loops run under OpenMP (timed CPU baseline); results are
 *           identical to threads==1.
 * Returns 0, or <0 on capacity/argument errors.
 */
int orc_make_tree(const orc_params *p, int32_t W, const int64_t *row_off,
                  const int32_t *x, int32_t S, int32_t F,
                  const float *gp, int32_t gid,
                  const int32_t *start, const int32_t *fea_index, int32_t n_fea,
                  const uint8_t *row_mask,
                  orc_tree *out, int32_t *node_of_row, int32_t *hist_miss,
                  int32_t threads, double *time_stats /* nullable, 4 doubles: hist, split, partition, total (s) */,
                  double *hist_cells /* nullable: sum over built nodes of rows*features */);

/* Tree.getLeafIndex on slot data with `<=` (J/data/gbdt/Tree.java:142-168,
 * J/optimizer/GBDTOptimizer.java:641-658): adds leaf value into float score[(row*K+gid)]. */
void orc_score_slots(const orc_tree *t, const int32_t *x, int64_t n, int32_t S,
                     float *score, int32_t K, int32_t gid);

/* Same walk on real values against real thresholds (isOriginTree=true). thr[nid]
 * replaces node_info for inner nodes. */
void orc_score_values(const orc_tree *t, const float *thr, const float *xv, int64_t n,
                      int32_t S, float *score, int32_t K, int32_t gid);

/* ... and on the int32 view of the values, with Constants.INT_MISSING_VALUE (0x7fc00000) taking
 * def_child[nid] (null: the left child) -- J/data/gbdt/Tree.java:142-168. */
void orc_score_value_bits(const orc_tree *t, const float *thr, const int32_t *def_child, const int32_t *xbits,
                          int64_t n, int32_t S, float *score, int32_t K, int32_t gid);

/* GBDTOptimizer.predictAndCalcLossGrad, numTreeInGroup==1 branch (:536-557) with
 * SigmoidFunction (J/loss/SigmoidFunction.java:39-93, zMax==0) or L2Function
 * (J/loss/L2Function.java:30-55).  loss_kind 0 = sigmoid, 1 = l2.
 * Returns sum_i w*loss; writes gp (nullable). */
double orc_loss_grad_scalar(int32_t loss_kind, const float *score, const float *init_score,
                            const float *y, const float *w, int64_t n, float *gp);

/* TreeRefiner.refine, exact data-parallel variant (J/optimizer/gbdt/TreeRefiner.java:104-123,209-255;
 * J/utils/PreciseQuantile.java:156-208): every leaf's value becomes learning_rate * weighted median of
 * `residual` over the leaf's rows (weight NULL: all 1).  loss_kind 3 of orc_loss_grad_scalar is L1Function. */
void orc_refine_leaves(orc_tree *t, const int32_t *node_of_row, const float *residual, const float *weight,
                       int64_t n, float learning_rate);

/* Softmax branch (:559-607, J/loss/SoftmaxFunction.java:44-113). y one-hot [n][K]. */
double orc_loss_grad_softmax(const float *score, const float *init_score, const float *y,
                             const float *w, int64_t n, int32_t K, float *gp, float *pred);

#ifdef __cplusplus
}
#endif
#endif
