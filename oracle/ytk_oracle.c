/*
 * ytk_oracle.c -- CPU restatement of ytk-learn's data-parallel GBDT tree builder.
 * TEST INFRASTRUCTURE ONLY (see ytk_oracle.h).  Citations: J/ =
 * /root/reference/src/main/java/com/fenbi/ytklearn/.
 *
 * Compile with -ffp-contract=off: Java never fuses a*b+c, so neither may we.
 */
#include "ytk_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

/* ------------------------------------------------------------------ A1 */

/* BinarySearch.findLastEqualOrUpper (J/utils/BinarySearch.java:33-56). */
static int32_t find_last_equal_or_upper(const float *arr, int32_t len, float target) {
    if (target > arr[len - 1]) return -1;
    int32_t l = 0, u = len - 1;
    while (l <= u) {
        int32_t mid = l + (u - l) / 2;
        if (target >= arr[mid]) l = mid + 1; else u = mid - 1;
    }
    if (u < 0) u = 0;
    if (arr[u] == target) return u;
    return (len - 1 < l) ? len - 1 : l;
}

/* Body of the sid loop in FeatureApprData.convertFeaVal2ApprFeaIndex
 * (J/data/gbdt/FeatureApprData.java:186-199).  `(c[i]+c[i-1])*0.5`: float add,
 * then double multiply (0.5 is a double literal); the compare is in double. */
int32_t orc_find_bin(const float *cuts, int32_t m, float v) {
    if (m == 1) return 0;
    int32_t idx = find_last_equal_or_upper(cuts, m, v);
    if (idx == -1) return m - 1;
    if (idx >= 1) {
        volatile float s = cuts[idx] + cuts[idx - 1];
        if ((double)v < (double)s * 0.5) idx -= 1;
    }
    return idx;
}

void orc_convert_to_bins(int32_t *x, int64_t n, int32_t S, int32_t F,
                         const float *cuts, const int32_t *start) {
    for (int32_t f = 0; f < F; f++) {                 /* fid outer, sid inner (:183-185) */
        const float *c = cuts + start[f];
        int32_t m = start[f + 1] - start[f];
        for (int64_t r = 0; r < n; r++) {
            float v;
            memcpy(&v, &x[r * S + f], 4);             /* NumConvertUtils.int2float */
            x[r * S + f] = orc_find_bin(c, m, v);
        }
    }
}

/* DataUtils.avgAssign (J/dataflow/DataUtils.java:193-205). */
void orc_avg_assign(int32_t amount, int32_t bins, int32_t *out) {
    int32_t avg = amount / bins, mod = amount % bins, fidx = 0;
    for (int32_t t = 0; t < bins; t++) {
        int32_t tnum = t < mod ? avg + 1 : avg;
        out[2 * t] = fidx;
        out[2 * t + 1] = fidx + tnum;
        fidx += tnum;
    }
}

/* J/param/gbdt/GBDTOptimizationParams.java:148-154. */
int32_t orc_normalize_max_leaf_cnt(int32_t max_depth, int32_t max_leaf_cnt) {
    if (max_depth > 0) {
        int32_t cap = (int32_t)pow(2.0, (double)max_depth);
        if (max_leaf_cnt < 0) return cap;
        return max_leaf_cnt < cap ? max_leaf_cnt : cap;
    }
    return max_leaf_cnt;
}

/* ------------------------------------------------------------ A4 gain math */

/* UpdateStrategy.thresholdL1 (J/optimizer/gbdt/UpdateStrategy.java:103-109). */
static double threshold_l1(double w, double lambda) {
    if (w > lambda) return w - lambda;
    if (w < -lambda) return w + lambda;
    return 0.0;
}

/* UpdateStrategy.calcNodeValue (:83-100). */
double orc_calc_node_value(const orc_params *p, double G, double H) {
    if (H < p->min_child_hessian_sum) return 0.0;
    double val;
    if (p->l1 == 0.0) val = -G / (H + p->l2);
    else val = -threshold_l1(G, p->l1) / (H + p->l2);
    if (p->max_abs_leaf_val > 0) {
        if (val > p->max_abs_leaf_val) val = p->max_abs_leaf_val;
        else if (val < -p->max_abs_leaf_val) val = -p->max_abs_leaf_val;
    }
    return val;
}

/* UpdateStrategy.calcGain (:64-80).  Math.pow(x,2) == x*x exactly. */
double orc_calc_gain(const orc_params *p, double G, double H) {
    if (H < p->min_child_hessian_sum) return 0.0;
    double gain;
    if (p->max_abs_leaf_val <= 0) {
        if (p->l1 == 0.0) {
            gain = (G * G) / (H + p->l2);
        } else {
            double t = threshold_l1(G, p->l1);
            gain = (t * t) / (H + p->l2);
        }
    } else {
        double v = orc_calc_node_value(p, G, H);
        gain = -2 * (G * v + 0.5 * (H + p->l2) * (v * v) + p->l1 * fabs(v));
    }
    return gain;
}

/* ------------------------------------------------------------------ A2 */

/* A "virtual" dense matrix for shapes whose dense int32 form does not fit in memory (BASELINE config 4:
 * 10 000 features at 1 % density): rows are held as CSR (entries of a row sorted by feature) plus one
 * default slot per feature, and x[row][fid] is looked up on the fly.  The loops below are still the
 * reference's dense loops (every row visits every sampled feature); only the storage of x differs.
 * Set once before orc_make_tree(x = NULL); read-only afterwards. */
static struct { const int64_t *row_ptr; const int32_t *fids; const int32_t *slots; const int32_t *default_slot; } g_csr;
void orc_set_virtual_csr(const int64_t *row_ptr, const int32_t *fids, const int32_t *slots, const int32_t *default_slot) {
    g_csr.row_ptr = row_ptr; g_csr.fids = fids; g_csr.slots = slots; g_csr.default_slot = default_slot;
}
static int32_t csr_slot(int64_t row, int32_t fid) {
    int64_t lo = g_csr.row_ptr[row], hi = g_csr.row_ptr[row + 1];
    const int64_t end = hi;
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if (g_csr.fids[mid] < fid) lo = mid + 1; else hi = mid;
    }
    return (lo < end && g_csr.fids[lo] == fid) ? g_csr.slots[lo] : g_csr.default_slot[fid];
}

/* Accumulation loop of HistogramBuilder.buildHist (J/data/gbdt/HistogramBuilder.java:69-90).
 * x == NULL: the virtual matrix above, rows offset by row_base. */
static void build_hist_rows(const int32_t *x, int32_t S, int64_t row_base, const float *gp, int32_t K, int32_t gid,
                            const int32_t *pos_idx, int64_t v1, int64_t v2,
                            const int32_t *fea_index, int32_t n_fea, const int32_t *start,
                            int32_t B, double *hist) {
    for (int64_t i = 0; i < 2 * (int64_t)B; i++) hist[i] = 0;
    for (int64_t i = v1; i < v2; i++) {
        if (pos_idx[2 * i] < 0) continue;                       /* not sampled */
        int64_t row = pos_idx[2 * i + 1];
        int64_t gi = (row * K + gid) * 2;
        float g = gp[gi], h = gp[gi + 1];
        if (x) {
            const int32_t *xr = x + row * S;
            for (int32_t k = 0; k < n_fea; k++) {
                int32_t fid = fea_index[k];
                int64_t b = ((int64_t)start[fid] + xr[fid]) * 2;
                hist[b] += g;                                   /* float widened to double */
                hist[b + 1] += h;
            }
        } else {
            const int64_t p0 = g_csr.row_ptr[row_base + row], p1 = g_csr.row_ptr[row_base + row + 1];
            int64_t q = p0;
            int32_t prev = -1;
            for (int32_t k = 0; k < n_fea; k++) {
                int32_t fid = fea_index[k];
                if (fid < prev) q = p0;                         /* feaIndex is sorted in the reference (:449) */
                prev = fid;
                while (q < p1 && g_csr.fids[q] < fid) q++;
                int32_t slot = (q < p1 && g_csr.fids[q] == fid) ? g_csr.slots[q] : g_csr.default_slot[fid];
                int64_t b = ((int64_t)start[fid] + slot) * 2;
                hist[b] += g;
                hist[b + 1] += h;
            }
        }
    }
}
void orc_build_hist(const int32_t *x, int32_t S, const float *gp, int32_t K, int32_t gid,
                    const int32_t *pos_idx, int64_t v1, int64_t v2,
                    const int32_t *fea_index, int32_t n_fea, const int32_t *start,
                    int32_t B, double *hist) {
    build_hist_rows(x, S, 0, gp, K, gid, pos_idx, v1, v2, fea_index, n_fea, start, B, hist);
}

/* ------------------------------------------------------------------ A5/A6 */

void orc_split_clear(orc_split *s) {                            /* SplitInfo.clear (:90-97) */
    s->loss_chg = -INFINITY; s->fid = -1; s->lo = -1; s->hi = -1;
}

/* SplitInfo.needReplace (J/data/gbdt/SplitInfo.java:99-104). */
int orc_split_need_replace(const orc_split *cur, float new_loss_chg, int32_t new_fid) {
    if (cur->fid <= new_fid) return new_loss_chg > cur->loss_chg;
    return new_loss_chg >= cur->loss_chg;
}

/* DataParallelTreeMaker.enumerateSplit (:598-637) + hasSample (:657-663). */
void orc_enumerate_split(const orc_params *p, const double *hist, const int32_t *start,
                         int32_t fid, double sumG, double sumH, float root_gain,
                         orc_split *best) {
    double lG = 0.0, lH = 0.0;                                  /* leftBranchGradStats.clear() */
    int32_t last = 0;
    int32_t nslot = start[fid + 1] - start[fid];
    const double *h = hist + 2 * (int64_t)start[fid];
    for (int32_t j = 0; j < nslot; j++) {
        double g = h[2 * j], hh = h[2 * j + 1];
        if (g == 0.0 && hh == 0.0) continue;                    /* hasSample */
        if (lH == 0.0) {                                        /* GradStats.empty(): sumHess == 0 */
            lG += g; lH += hh; last = j;
        } else {
            if (lH >= p->min_child_hessian_sum) {
                double rG = sumG - lG, rH = sumH - lH;          /* setSubstract */
                if (rH >= p->min_child_hessian_sum) {
                    float loss_chg = (float)(orc_calc_gain(p, lG, lH) + orc_calc_gain(p, rG, rH)
                                             - (double)root_gain);
                    if (orc_split_need_replace(best, loss_chg, fid)) {  /* SplitInfo.update (:130-140) */
                        best->loss_chg = loss_chg; best->fid = fid; best->lo = last; best->hi = j;
                    }
                }
            }
            lG += g; lH += hh; last = j;
        }
    }
}

/* ------------------------------------------------------------------ A7 */

/* SamplePositionData.sortByNodeIndex (J/data/gbdt/SamplePositionData.java:136-165). */
static int64_t sort_by_node_index(int32_t *pi, int64_t v1, int64_t v2, int32_t left, int32_t right) {
    int64_t ps = v1, pe = v2 - 1, left_cnt = 0;
    while (ps <= pe) {
        while (ps <= pe && pi[ps << 1] == left) { ps++; left_cnt++; }
        while (ps <= pe && pi[pe << 1] == right) pe--;
        if (ps < pe && pi[ps << 1] > pi[pe << 1]) {
            int32_t t = pi[2 * ps]; pi[2 * ps] = pi[2 * pe]; pi[2 * pe] = t;
            t = pi[2 * ps + 1]; pi[2 * ps + 1] = pi[2 * pe + 1]; pi[2 * pe + 1] = t;
            ps++; pe--; left_cnt++;
        }
    }
    return left_cnt;
}

/* SamplePositionData.resetPosition(trainData, nid, tree) (:115-133): strict `<` on a
 * float cond against an int slot. */
static int64_t reset_position_rows(int32_t *pos_idx, int64_t v1, int64_t v2, const int32_t *x,
                                   int32_t S, int64_t row_base, int32_t fid, float cond, int32_t left, int32_t right) {
    for (int64_t i = v1; i < v2; i++) {
        int64_t row = pos_idx[2 * i + 1];
        int32_t slot = x ? x[row * S + fid] : csr_slot(row_base + row, fid);
        if ((float)slot < cond) pos_idx[2 * i] = left;
        else pos_idx[2 * i] = right;
    }
    return sort_by_node_index(pos_idx, v1, v2, left, right);
}
int64_t orc_reset_position(int32_t *pos_idx, int64_t v1, int64_t v2, const int32_t *x,
                           int32_t S, int32_t fid, float cond, int32_t left, int32_t right) {
    return reset_position_rows(pos_idx, v1, v2, x, S, 0, fid, cond, left, right);
}

/* ------------------------------------------------------------------ A3' pool */

/* HistogramPool (J/data/gbdt/HistogramPool.java:190-267).  Items keep their buffers
 * when recycled, exactly as the Java arrays do, so a caller holding a buffer pointer
 * across an eviction sees the new contents (as the reference would). */
typedef struct { int32_t nid; double *buf; } pool_item;
typedef struct {
    int32_t capacity, cur_size, n_used, n_avail;
    int64_t len;
    pool_item **used;    /* allocation order, oldest first  */
    pool_item **avail;   /* FIFO                             */
    pool_item *items;
} hist_pool;

static int pool_init(hist_pool *hp, int32_t capacity, int64_t len) {
    memset(hp, 0, sizeof(*hp));
    if (capacity < 1) return -1;
    hp->capacity = capacity; hp->len = len;
    hp->items = (pool_item *)calloc((size_t)capacity, sizeof(pool_item));
    hp->used = (pool_item **)calloc((size_t)capacity, sizeof(pool_item *));
    hp->avail = (pool_item **)calloc((size_t)capacity, sizeof(pool_item *));
    return (hp->items && hp->used && hp->avail) ? 0 : -1;
}
static void pool_free(hist_pool *hp) {
    if (hp->items) for (int32_t i = 0; i < hp->cur_size; i++) free(hp->items[i].buf);
    free(hp->items); free(hp->used); free(hp->avail);
}
static double *pool_get(hist_pool *hp, int32_t nid) {           /* getHist (:238-245) */
    for (int32_t i = 0; i < hp->n_used; i++) if (hp->used[i]->nid == nid) return hp->used[i]->buf;
    return NULL;
}
static double *pool_alloc(hist_pool *hp, int32_t nid) {         /* allocHist (:209-236) */
    pool_item *it;
    if (hp->n_avail > 0) {
        it = hp->avail[0];
        memmove(hp->avail, hp->avail + 1, sizeof(pool_item *) * (size_t)(hp->n_avail - 1));
        hp->n_avail--;
    } else if (hp->cur_size < hp->capacity) {
        it = &hp->items[hp->cur_size++];
        it->buf = (double *)calloc((size_t)hp->len, sizeof(double));
        if (!it->buf) return NULL;
    } else {
        it = hp->used[0];                                       /* evict oldest */
        memmove(hp->used, hp->used + 1, sizeof(pool_item *) * (size_t)(hp->n_used - 1));
        hp->n_used--;
    }
    it->nid = nid;
    hp->used[hp->n_used++] = it;
    return it->buf;
}
static double *pool_set(hist_pool *hp, int32_t nid, const double *h) { /* setHist (:247-254) */
    double *b = pool_alloc(hp, nid);
    if (b) memcpy(b, h, sizeof(double) * (size_t)hp->len);
    return b;
}
static void pool_release(hist_pool *hp, int32_t nid) {          /* releaseHist (:256-267) */
    for (int32_t i = 0; i < hp->n_used; i++) {
        if (hp->used[i]->nid == nid) {
            pool_item *it = hp->used[i];
            memmove(hp->used + i, hp->used + i + 1, sizeof(pool_item *) * (size_t)(hp->n_used - 1 - i));
            hp->n_used--;
            hp->avail[hp->n_avail++] = it;
            return;
        }
    }
}

/* ------------------------------------------------------------------ make() */

typedef struct {
    double sumG, sumH;
    float root_gain, node_value;
    int64_t sample_cnt;
    orc_split best;
} node_stats;

typedef struct { int32_t nid; float loss_chg; int32_t depth, seq; } expand_node;

/* java.lang.Float.compare */
static int java_float_compare(float a, float b) {
    if (a < b) return -1;
    if (a > b) return 1;
    int32_t ia, ib;
    if (a != a) ia = 0x7fc00000; else memcpy(&ia, &a, 4);
    if (b != b) ib = 0x7fc00000; else memcpy(&ib, &b, 4);
    return ia == ib ? 0 : (ia < ib ? -1 : 1);
}
/* lossChgComparator / sequenceComparator (DataParallelTreeMaker.java:105-115): <0 ⇒ a first */
static int expand_cmp(const expand_node *a, const expand_node *b, int loss_wise) {
    if (loss_wise) {
        int r = java_float_compare(b->loss_chg, a->loss_chg);
        if (r != 0) return r;
    }
    return (a->seq > b->seq) - (a->seq < b->seq);
}

typedef struct {
    const orc_params *p;
    int32_t W, S, F, K, gid, B, n_fea, loss_wise;
    const int64_t *row_off;
    const int32_t *x;
    const float *gp;
    const int32_t *start, *fea_index;
    int32_t *fea_assign;          /* 2*W: [from,to) features of worker w           */
    int32_t *first_fea;           /* firstSampledFeaIndexInThread per worker, -1   */
    int32_t **pos;                /* per worker (nid,row) pairs                    */
    int64_t *iv;                  /* intervals: iv[(nid*W + w)*2 + {0,1}]          */
    double **local;               /* per worker local histogram (2B)               */
    double *small_sum, *large_sum;/* smallLeafGradSum / largeLeafGradSum (2B)      */
    hist_pool pool;
    node_stats *ns;
    orc_tree *t;
    int32_t threads, hist_miss;
    double t_hist, t_split, t_part, cells;
} maker;

static int32_t alloc_node(orc_tree *t) {                        /* Tree.AllocTreeNode (:58-64) */
    if (t->n_nodes >= t->capacity) return -1;
    int32_t nid = t->n_nodes++;
    t->parent[nid] = -1; t->left[nid] = -1; t->right[nid] = -1;
    t->fid[nid] = 0; t->lo[nid] = 0; t->hi[nid] = 0; t->node_info[nid] = 0.0f;
    t->loss_chg[nid] = 0.0f; t->hess_sum[nid] = 0.0f; t->sample_cnt[nid] = 0;
    return nid;
}

static void ns_clear(node_stats *s) {
    s->sumG = s->sumH = 0.0; s->root_gain = 0.0f; s->node_value = 0.0f; s->sample_cnt = 0;
    orc_split_clear(&s->best);
}

/* syncGlobalGradSum (:478-483): every worker builds its local histogram, the
 * reduce-scatter sums them (rank order), the result goes into the pool. */
static double *sync_global_grad_sum(maker *m, int32_t nid, double *scratch) {
    double t0 = now_s();
    int64_t rows = 0;
#pragma omp parallel for num_threads(m->threads) if (m->threads > 1) schedule(static, 1)
    for (int32_t w = 0; w < m->W; w++) {
        int64_t v1 = m->iv[((int64_t)nid * m->W + w) * 2], v2 = m->iv[((int64_t)nid * m->W + w) * 2 + 1];
        build_hist_rows(m->x ? m->x + m->row_off[w] * m->S : NULL, m->S, m->row_off[w], m->gp + m->row_off[w] * m->K * 2,
                        m->K, m->gid, m->pos[w], v1, v2, m->fea_index, m->n_fea, m->start, m->B, m->local[w]);
    }
    for (int32_t w = 0; w < m->W; w++)
        rows += m->iv[((int64_t)nid * m->W + w) * 2 + 1] - m->iv[((int64_t)nid * m->W + w) * 2];
    m->cells += (double)rows * (double)m->n_fea;
    /* reduceScatterArray(SUM) (HistogramBuilder.java:95): owner segments, summed in rank order */
    int64_t len = 2 * (int64_t)m->B;
    if (m->W == 1) {
        memcpy(scratch, m->local[0], sizeof(double) * (size_t)len);
    } else {
#pragma omp parallel for num_threads(m->threads) if (m->threads > 1)
        for (int64_t i = 0; i < len; i++) {
            double s = m->local[0][i];
            for (int32_t w = 1; w < m->W; w++) s += m->local[w][i];
            scratch[i] = s;
        }
    }
    m->t_hist += now_s() - t0;
    return pool_set(&m->pool, nid, scratch);
}

/* updateTreeMakerNodeStats(nid, hist, needComm=true) (:543-573). */
static void update_node_stats_from_hist(maker *m, int32_t nid, const double *hist) {
    node_stats *s = &m->ns[nid];
    double G = -INFINITY, H = -INFINITY;
    for (int32_t w = 0; w < m->W; w++) {
        int32_t f = m->first_fea[w];
        if (f == -1) continue;
        double g = 0.0, h = 0.0;                                /* this worker's cleared sumGradStats */
        for (int32_t i = m->start[f]; i < m->start[f + 1]; i++) { g += hist[2 * i]; h += hist[2 * i + 1]; }
        if (g > G) G = g;                                       /* allreduce MAX, element-wise (:560) */
        if (h > H) H = h;
    }
    s->sumG = G; s->sumH = H;
    s->root_gain = (float)orc_calc_gain(m->p, G, H);
    s->node_value = (float)orc_calc_node_value(m->p, G, H);
}

/* findBestSplit(nid, hist) (:575-595) + syncBestSplit (:640-653). */
static void find_best_split(maker *m, int32_t nid, const double *hist) {
    node_stats *s = &m->ns[nid];
    /* UpdateStrategy.canSplit (:50-53) */
    if (!(s->sumH >= m->p->min_child_hessian_sum * 2.0 &&
          s->sample_cnt >= (int64_t)m->p->min_split_samples))
        return;
    orc_split merged; orc_split_clear(&merged);
    for (int32_t w = 0; w < m->W; w++) {
        orc_split best; orc_split_clear(&best);
        for (int32_t k = 0; k < m->n_fea; k++) {
            int32_t fid = m->fea_index[k];
            if (fid >= m->fea_assign[2 * w] && fid < m->fea_assign[2 * w + 1])
                orc_enumerate_split(m->p, hist, m->start, fid, s->sumG, s->sumH, s->root_gain, &best);
        }
        if (w == 0) merged = best;
        else if (orc_split_need_replace(&merged, best.loss_chg, best.fid)) merged = best;
    }
    s->best = merged;
    if ((double)merged.loss_chg > m->p->min_split_loss) {       /* TreeNode.setSplit (:144-149) */
        m->t->fid[nid] = merged.fid; m->t->lo[nid] = merged.lo; m->t->hi[nid] = merged.hi;
        m->t->node_info[nid] = 0.5f * (float)(merged.lo + merged.hi);
    }
}

static void set_leaf(maker *m, int32_t nid) {                   /* :253, TreeNode.setLeaf */
    m->t->node_info[nid] = m->ns[nid].node_value * m->p->learning_rate;
    m->t->left[nid] = -1; m->t->right[nid] = -1;
}

int orc_make_tree(const orc_params *p, int32_t W, const int64_t *row_off,
                  const int32_t *x, int32_t S, int32_t F,
                  const float *gp, int32_t gid,
                  const int32_t *start, const int32_t *fea_index, int32_t n_fea,
                  const uint8_t *row_mask,
                  orc_tree *out, int32_t *node_of_row, int32_t *hist_miss,
                  int32_t threads, double *time_stats, double *hist_cells) {
    if (W < 1 || n_fea < 1 || out->capacity < 1) return -1;
    double t_total0 = now_s();
    maker m; memset(&m, 0, sizeof(m));
    m.p = p; m.W = W; m.S = S; m.F = F; m.K = p->class_num; m.gid = gid; m.B = start[F];
    m.n_fea = n_fea; m.loss_wise = p->grow_policy == 1; m.row_off = row_off; m.x = x; m.gp = gp;
    m.start = start; m.fea_index = fea_index; m.t = out; m.threads = threads < 1 ? 1 : threads;
    int rc = 0;
    int32_t max_nodes = out->capacity;
    int64_t len = 2 * (int64_t)m.B;

    /* pool capacity (:193-203); HistogramPool.UNIT = 64, Constants.MB = 1<<20 */
    int32_t capacity;
    if (p->histogram_pool_capacity > 0) {
        capacity = (int32_t)(p->histogram_pool_capacity * 1048576.0 / 64 / m.B);
        if (capacity < 2) capacity = 2;
        if (capacity > p->max_leaf_cnt) capacity = p->max_leaf_cnt;
    } else {
        capacity = p->max_leaf_cnt;
    }
    if (pool_init(&m.pool, capacity, len) != 0) return -2;

    m.fea_assign = (int32_t *)malloc(sizeof(int32_t) * 2 * (size_t)W);
    m.first_fea = (int32_t *)malloc(sizeof(int32_t) * (size_t)W);
    m.pos = (int32_t **)calloc((size_t)W, sizeof(int32_t *));
    m.local = (double **)calloc((size_t)W, sizeof(double *));
    m.iv = (int64_t *)calloc((size_t)max_nodes * (size_t)W * 2, sizeof(int64_t));
    m.small_sum = (double *)calloc((size_t)len, sizeof(double));
    m.large_sum = (double *)calloc((size_t)len, sizeof(double));
    m.ns = (node_stats *)calloc((size_t)max_nodes, sizeof(node_stats));
    expand_node *queue = (expand_node *)calloc((size_t)max_nodes, sizeof(expand_node));
    int32_t qn = 0;

    /* GBDTDataFlow.handleOtherTrainInfo (J/dataflow/GBDTDataFlow.java:252-263): xColRange */
    orc_avg_assign(F, W, m.fea_assign);
    /* initAssistData (:395-470): positions, row subsample, firstSampledFeaIndexInThread */
    for (int32_t w = 0; w < W; w++) {
        int64_t n = row_off[w + 1] - row_off[w];
        m.pos[w] = (int32_t *)malloc(sizeof(int32_t) * 2 * (size_t)(n > 0 ? n : 1));
        m.local[w] = (double *)calloc((size_t)len, sizeof(double));
        int64_t v1 = 0;
        for (int64_t i = 0; i < n; i++) {                       /* SamplePositionData.clear (:55-62) */
            m.pos[w][2 * i] = 0; m.pos[w][2 * i + 1] = (int32_t)i;
        }
        if (row_mask) {
            for (int64_t i = 0; i < n; i++) if (!row_mask[row_off[w] + i]) m.pos[w][2 * i] = -1;
            v1 = sort_by_node_index(m.pos[w], 0, n, -1, 0);     /* resetPositionAfterSample (:84-87) */
        }
        m.iv[(0 * (int64_t)W + w) * 2] = v1; m.iv[(0 * (int64_t)W + w) * 2 + 1] = n;
        m.first_fea[w] = -1;
        if (m.fea_assign[2 * w] < m.fea_assign[2 * w + 1]) {
            for (int32_t k = 0; k < n_fea; k++) {
                if (fea_index[k] >= m.fea_assign[2 * w] && fea_index[k] < m.fea_assign[2 * w + 1]) {
                    m.first_fea[w] = fea_index[k]; break;
                }
            }
        }
    }

    out->n_nodes = 0;
    int32_t root = alloc_node(out);                             /* tree.init() */
    int32_t seq = 0;
    /* initTreeMakerNodeStats(0) (:510-520) */
    ns_clear(&m.ns[root]);
    for (int32_t w = 0; w < W; w++) m.ns[root].sample_cnt += m.iv[w * 2 + 1] - m.iv[w * 2];
    /* findBestSplit(root) (:302-318) */
    {
        double *h = sync_global_grad_sum(&m, root, m.small_sum);
        if (!h) { rc = -3; goto done; }
        double t0 = now_s();
        update_node_stats_from_hist(&m, root, h);
        find_best_split(&m, root, h);
        m.t_split += now_s() - t0;
    }
    queue[qn++] = (expand_node){root, m.ns[root].best.loss_chg, 0, seq++};

    int32_t num_leaf = 1;
    while (qn > 0) {
        int32_t bi = 0;                                         /* PriorityQueue.poll() */
        for (int32_t i = 1; i < qn; i++) if (expand_cmp(&queue[i], &queue[bi], m.loss_wise) < 0) bi = i;
        expand_node e = queue[bi];
        queue[bi] = queue[--qn];
        int32_t nid = e.nid;

        if ((double)e.loss_chg <= p->min_split_loss
            || (p->max_depth >= 0 && p->max_depth == e.depth)
            || (p->max_leaf_cnt > 0 && p->max_leaf_cnt == num_leaf)
            || (p->min_split_samples > 0 && m.ns[nid].sample_cnt < p->min_split_samples)) {
            set_leaf(&m, nid);
            continue;
        }
        int32_t l = alloc_node(out), r = alloc_node(out);       /* tree.addChilds (:66-73) */
        if (l < 0 || r < 0) { rc = -4; goto done; }
        out->left[nid] = l; out->right[nid] = r; out->parent[l] = nid; out->parent[r] = nid;
        num_leaf++;

        /* position.resetPosition (SamplePositionData.java:115-133) */
        double t0 = now_s();
#pragma omp parallel for num_threads(m.threads) if (m.threads > 1) schedule(static, 1)
        for (int32_t w = 0; w < W; w++) {
            int64_t v1 = m.iv[((int64_t)nid * W + w) * 2], v2 = m.iv[((int64_t)nid * W + w) * 2 + 1];
            int64_t lc = reset_position_rows(m.pos[w], v1, v2, x ? x + row_off[w] * S : NULL, S, row_off[w], out->fid[nid],
                                             out->node_info[nid], l, r);
            m.iv[((int64_t)l * W + w) * 2] = v1; m.iv[((int64_t)l * W + w) * 2 + 1] = v1 + lc;
            m.iv[((int64_t)r * W + w) * 2] = v1 + lc; m.iv[((int64_t)r * W + w) * 2 + 1] = v2;
        }
        m.t_part += now_s() - t0;

        /* initTreeMakerNodeStats(l, r) (:522-541) */
        ns_clear(&m.ns[l]); ns_clear(&m.ns[r]);
        for (int32_t w = 0; w < W; w++) {
            m.ns[l].sample_cnt += m.iv[((int64_t)l * W + w) * 2 + 1] - m.iv[((int64_t)l * W + w) * 2];
            m.ns[r].sample_cnt += m.iv[((int64_t)r * W + w) * 2 + 1] - m.iv[((int64_t)r * W + w) * 2];
        }
        /* smaller child: strictly fewer rows on the left, else the right (:344-350, :493-499) */
        int32_t small = (m.ns[l].sample_cnt < m.ns[r].sample_cnt) ? l : r;
        int32_t large = (small == l) ? r : l;

        if ((p->max_depth >= 0 && p->max_depth == e.depth + 1)
            || (p->max_leaf_cnt > 0 && p->max_leaf_cnt == num_leaf)
            || (p->min_split_samples > 0 && m.ns[l].sample_cnt < p->min_split_samples
                && m.ns[r].sample_cnt < p->min_split_samples)) {
            /* computeLeafStats (:341-385) */
            t0 = now_s();
            double tot[2] = {0, 0};
            for (int32_t w = 0; w < W; w++) {
                double g = 0, h = 0;
                int64_t v1 = m.iv[((int64_t)small * W + w) * 2], v2 = m.iv[((int64_t)small * W + w) * 2 + 1];
                const float *gpw = gp + row_off[w] * m.K * 2;
                for (int64_t i = v1; i < v2; i++) {
                    if (m.pos[w][2 * i] < 0) continue;
                    int64_t gi = ((int64_t)m.pos[w][2 * i + 1] * m.K + gid) * 2;
                    g += gpw[gi]; h += gpw[gi + 1];
                }
                tot[0] += g; tot[1] += h;                       /* allreduceArray(SUM) (:375) */
            }
            m.t_hist += now_s() - t0;
            node_stats *ss = &m.ns[small], *ls = &m.ns[large], *ps = &m.ns[nid];
            ss->sumG = tot[0]; ss->sumH = tot[1];
            ss->root_gain = (float)orc_calc_gain(p, ss->sumG, ss->sumH);
            ss->node_value = (float)orc_calc_node_value(p, ss->sumG, ss->sumH);
            ls->sumG = ps->sumG - ss->sumG; ls->sumH = ps->sumH - ss->sumH;
            ls->root_gain = (float)orc_calc_gain(p, ls->sumG, ls->sumH);
            ls->node_value = (float)orc_calc_node_value(p, ls->sumG, ls->sumH);
            set_leaf(&m, l); set_leaf(&m, r);
        } else {
            /* findBestSplit(parent, l, r) (:320-339) + buildHist(parent,l,r) (:489-508) */
            double *parent_h = pool_get(&m.pool, nid);
            double *small_h = sync_global_grad_sum(&m, small, m.small_sum);
            double *large_h;
            if (!small_h) { rc = -3; goto done; }
            if (parent_h) {
                t0 = now_s();
                /* HistogramBuilder.substract (:101-107) over every owner's segment */
                for (int64_t i = 0; i < len; i++) m.large_sum[i] = parent_h[i] - small_h[i];
                pool_release(&m.pool, nid);
                large_h = pool_set(&m.pool, large, m.large_sum);
                m.t_hist += now_s() - t0;
            } else {
                large_h = sync_global_grad_sum(&m, large, m.large_sum);
                m.hist_miss++;
            }
            if (!large_h) { rc = -3; goto done; }
            t0 = now_s();
            update_node_stats_from_hist(&m, small, small_h);
            update_node_stats_from_hist(&m, large, large_h);
            find_best_split(&m, small, small_h);
            find_best_split(&m, large, large_h);
            m.t_split += now_s() - t0;
            queue[qn++] = (expand_node){l, m.ns[l].best.loss_chg, e.depth + 1, seq++};
            queue[qn++] = (expand_node){r, m.ns[r].best.loss_chg, e.depth + 1, seq++};
        }
    }

    /* updateTreeNodeStat (:387-393) */
    for (int32_t nid = 0; nid < out->n_nodes; nid++) {
        out->loss_chg[nid] = m.ns[nid].best.loss_chg;
        out->hess_sum[nid] = (float)m.ns[nid].sumH;
        out->sample_cnt[nid] = m.ns[nid].sample_cnt;
    }
    if (node_of_row) {
        for (int32_t w = 0; w < W; w++) {
            int64_t n = row_off[w + 1] - row_off[w];
            for (int64_t i = 0; i < n; i++)
                node_of_row[row_off[w] + m.pos[w][2 * i + 1]] = m.pos[w][2 * i];
        }
    }
done:
    if (hist_miss) *hist_miss = m.hist_miss;
    if (hist_cells) *hist_cells = m.cells;
    if (time_stats) {
        time_stats[0] = m.t_hist; time_stats[1] = m.t_split; time_stats[2] = m.t_part;
        time_stats[3] = now_s() - t_total0;
    }
    for (int32_t w = 0; w < W; w++) { if (m.pos) free(m.pos[w]); if (m.local) free(m.local[w]); }
    free(m.pos); free(m.local); free(m.iv); free(m.small_sum); free(m.large_sum); free(m.ns);
    free(m.fea_assign); free(m.first_fea); free(queue);
    pool_free(&m.pool);
    return rc;
}

/* ------------------------------------------------------------------ scoring / gradients */

void orc_score_slots(const orc_tree *t, const int32_t *x, int64_t n, int32_t S,
                     float *score, int32_t K, int32_t gid) {
    for (int64_t r = 0; r < n; r++) {
        int32_t pid = 0;
        while (t->left[pid] != -1) {                            /* Tree.getLeafIndex (:142-154) */
            float v = (float)x[r * S + t->fid[pid]];
            pid = (v <= t->node_info[pid]) ? t->left[pid] : t->right[pid];   /* getNext (:156-168) */
        }
        score[r * K + gid] += t->node_info[pid];                /* accPred += curPred (float) */
    }
}

void orc_score_values(const orc_tree *t, const float *thr, const float *xv, int64_t n,
                      int32_t S, float *score, int32_t K, int32_t gid) {
    for (int64_t r = 0; r < n; r++) {
        int32_t pid = 0;
        while (t->left[pid] != -1) {
            float v = xv[r * S + t->fid[pid]];
            pid = (v <= thr[pid]) ? t->left[pid] : t->right[pid];
        }
        score[r * K + gid] += t->node_info[pid];
    }
}

/* The same walk on the int32 view of the values (GBDTCoreData.x): a value equal to
 * Constants.INT_MISSING_VALUE = floatToRawIntBits(Float.NaN) = 0x7fc00000 takes the node's default
 * child (Tree.getLeafIndex :147-148, getNext :156-160; TreeNode.getDefualtChild). */
void orc_score_value_bits(const orc_tree *t, const float *thr, const int32_t *def_child, const int32_t *xbits,
                          int64_t n, int32_t S, float *score, int32_t K, int32_t gid) {
    for (int64_t r = 0; r < n; r++) {
        int32_t pid = 0;
        while (t->left[pid] != -1) {
            int32_t bits = xbits[r * S + t->fid[pid]];
            if (bits == 0x7fc00000) {
                pid = def_child ? def_child[pid] : t->left[pid];
            } else {
                float v;
                memcpy(&v, &bits, 4);
                pid = (v <= thr[pid]) ? t->left[pid] : t->right[pid];
            }
        }
        score[r * K + gid] += t->node_info[pid];
    }
}

double orc_loss_grad_scalar(int32_t loss_kind, const float *score, const float *init_score,
                            const float *y, const float *w, int64_t n, float *gp) {
    double loss = 0.0;
    for (int64_t i = 0; i < n; i++) {
        double wei = w[i];
        double s = (double)(float)(score[i] + init_score[i]);   /* float + float, then widened (:545) */
        double label = y[i];
        double l, pred;
        if (loss_kind == 0) {                                   /* SigmoidFunction.loss / predict */
            if (s >= 0.0) l = log(1 + exp(-s)) + s * (1.0 - label);
            else l = log(1 + exp(s)) - s * label;
            if (s >= 0.0) pred = 1.0 / (1.0 + exp(-s));
            else { double ez = exp(s); pred = ez / (1.0 + ez); }
        } else if (loss_kind == 3) {                            /* L1Function (J/loss/L1Function.java:31-54) */
            l = fabs(label - s);
            pred = s;
        } else {                                                /* L2Function */
            l = 0.5 * (label - s) * (label - s);
            pred = s;
        }
        loss += wei * l;
        float pf = (float)pred;                                 /* data.predict = (float) obj.predict */
        if (gp) {
            double p = pf, d0, d1;
            if (loss_kind == 0) { d0 = p - label; d1 = p * (1.0 - p); }   /* getDerivativeFast */
            else if (loss_kind == 3) { d0 = p - label; d0 = d0 > 0 ? 1.0 : (d0 < 0 ? -1.0 : d0); d1 = 1.0; }   /* Math.signum */
            else { d0 = p - label; d1 = 1.0; }                  /* ILossFunction default: first/second derivative */
            gp[2 * i] = (float)(d0 * wei);
            gp[2 * i + 1] = (float)(d1 * wei);
        }
    }
    return loss;
}

double orc_loss_grad_softmax(const float *score, const float *init_score, const float *y,
                             const float *w, int64_t n, int32_t K, float *gp, float *pred) {
    double loss = 0.0;
    double *ls = (double *)malloc(sizeof(double) * (size_t)K * 2);
    double *lp = ls + K;
    for (int64_t i = 0; i < n; i++) {
        double wei = w[i];
        for (int32_t k = 0; k < K; k++) ls[k] = (double)(float)(score[i * K + k] + init_score[i * K + k]);
        double mx = ls[K - 1];                                  /* SoftmaxFunction.loss (:44-63) */
        for (int32_t k = 0; k < K - 1; k++) if (ls[k] > mx) mx = ls[k];
        double esum = 0.0, sig = 0.0;
        for (int32_t k = 0; k < K; k++) { double nw = ls[k] - mx; sig += nw * (double)y[i * K + k]; esum += exp(nw); }
        loss += wei * (log(esum) - sig);
        esum = 0.0;                                             /* predict (:65-86) */
        for (int32_t k = 0; k < K; k++) { lp[k] = exp(ls[k] - mx); esum += lp[k]; }
        double inv = 1.0 / esum;
        for (int32_t k = 0; k < K; k++) lp[k] = lp[k] * inv;
        for (int32_t k = 0; k < K; k++) {
            if (pred) pred[i * K + k] = (float)lp[k];
            if (gp) {                                           /* getDerivativeFast (:108-113) on double pred */
                double d0 = lp[k] - (double)y[i * K + k];
                double d1 = 2 * (lp[k] * (1 - lp[k]));
                gp[(i * K + k) * 2] = (float)(d0 * wei);
                gp[(i * K + k) * 2 + 1] = (float)(d1 * wei);
            }
        }
    }
    free(ls);
    return loss;
}

/* ------------------------------------------------------------------ TreeRefiner (l1) */

typedef struct { float v, w; } vw_pair;
static int vw_cmp(const void *a, const void *b) {
    float x = ((const vw_pair *)a)->v, y = ((const vw_pair *)b)->v;
    return x < y ? -1 : (x > y ? 1 : 0);
}

/* TreeRefiner.refine, exact data-parallel variant (J/optimizer/gbdt/TreeRefiner.java:104-123,209-255) with
 * PreciseQuantile's weighted median (J/utils/PreciseQuantile.java:156-208: all workers' (value, weight)
 * pairs of a leaf concatenated, sorted by value, weights accumulated in double in that order, the first
 * value whose accumulated weight reaches half the total; an empty leaf gives QUANTILE_MISSGING = +inf):
 * leaf value = median(label - score) * learningRate.  residual/weight are per row; node_of_row as
 * orc_make_tree returns it (-1: the row was not sampled). */
void orc_refine_leaves(orc_tree *t, const int32_t *node_of_row, const float *residual, const float *weight,
                       int64_t n, float learning_rate) {
    int64_t *cnt = (int64_t *)calloc((size_t)t->n_nodes + 1, sizeof(int64_t));
    for (int64_t i = 0; i < n; i++) if (node_of_row[i] >= 0) cnt[node_of_row[i] + 1]++;
    for (int32_t k = 0; k < t->n_nodes; k++) cnt[k + 1] += cnt[k];
    vw_pair *buf = (vw_pair *)malloc(sizeof(vw_pair) * (size_t)(cnt[t->n_nodes] > 0 ? cnt[t->n_nodes] : 1));
    int64_t *cur = (int64_t *)malloc(sizeof(int64_t) * ((size_t)t->n_nodes + 1));
    memcpy(cur, cnt, sizeof(int64_t) * ((size_t)t->n_nodes + 1));
    for (int64_t i = 0; i < n; i++) {
        if (node_of_row[i] < 0) continue;
        vw_pair p = {residual[i], weight ? weight[i] : 1.0f};
        buf[cur[node_of_row[i]]++] = p;
    }
    for (int32_t nid = 0; nid < t->n_nodes; nid++) {
        if (t->left[nid] != -1) continue;
        vw_pair *seg = buf + cnt[nid];
        int64_t m = cnt[nid + 1] - cnt[nid];
        float med = INFINITY;
        if (m > 0) {
            qsort(seg, (size_t)m, sizeof(vw_pair), vw_cmp);
            double total = 0.0;
            for (int64_t i = 0; i < m; i++) total += seg[i].w;
            double target = total * 0.5, acc = 0.0;
            med = seg[m - 1].v;
            for (int64_t i = 0; i < m; i++) {
                acc += seg[i].w;
                if (acc >= target) { med = seg[i].v; break; }
            }
        }
        t->node_info[nid] = med * learning_rate;
    }
    free(buf); free(cur); free(cnt);
}
