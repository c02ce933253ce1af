/*
 * GpuTreeMaker -- the B200 tree maker behind ytk-learn's ITreeMaker seam
 * (optimizer/gbdt/ITreeMaker.java:34-39), a drop-in for DataParallelTreeMaker
 * (optimizer/gbdt/DataParallelTreeMaker.java:49-664): same constructor arguments, same make(groupId)
 * contract.  It binds libytkgbdt.so (include/ytk_gbdt.h) through jni/ytk_jni.cpp.
 *
 * Selected in GBDTOptimizer.initDataForTrain (optimizer/GBDTOptimizer.java:315-321):
 *     } else if (Boolean.getBoolean("ytk.gbdt.gpu")) {
 *         treeMaker = new GpuTreeMaker(comm, commonParams.optimizationParams, trainData, position, feaApprData);
 *     } else { treeMaker = new DataParallelTreeMaker(...); }
 * with thread_num = number of GPUs: worker thread t drives device t (worker/TrainWorker.java:167-200).
 *
 * This image has no JDK; the file is written against the reference's sources and is compiled by a
 * maintainer inside the ytk-learn tree (package path optimizer/gbdt/).
 */
package com.fenbi.ytklearn.optimizer.gbdt;

import com.fenbi.mp4j.comm.ThreadCommSlave;
import com.fenbi.mp4j.exception.Mp4jException;
import com.fenbi.mp4j.operand.Operands;
import com.fenbi.mp4j.operator.Operators;
import com.fenbi.ytklearn.data.Tuple;
import com.fenbi.ytklearn.data.gbdt.FeatureApprData;
import com.fenbi.ytklearn.data.gbdt.SamplePositionData;
import com.fenbi.ytklearn.data.gbdt.TimeStats;
import com.fenbi.ytklearn.data.gbdt.Tree;
import com.fenbi.ytklearn.data.gbdt.TreeNodeStat;
import com.fenbi.ytklearn.dataflow.GBDTCoreData;
import com.fenbi.ytklearn.param.gbdt.GBDTOptimizationParams;
import com.fenbi.ytklearn.data.gbdt.TreeGrowPolicy;
import com.fenbi.ytklearn.utils.RandomUtils;

import java.util.Arrays;
import java.util.List;
import java.util.Random;

public class GpuTreeMaker implements ITreeMaker, AutoCloseable {
    static { System.loadLibrary("ytkgbdt_jni"); }           // libytkgbdt_jni.so, which links libytkgbdt.so

    private final ThreadCommSlave comm;
    private final GBDTOptimizationParams params;
    private final GBDTCoreData trainData;
    private final SamplePositionData position;
    private final int sampleNum, featureDim;
    private long handle;                                    // ytk_gbdt*
    // TimeStats keeps whole milliseconds (long); a tree takes a few ms here, so fractions are summed
    // in doubles and rounded when the trainer asks for the totals.
    private final double[] totalMs = new double[8];
    private final int[] nodeOfRow;

    /** Same arguments as DataParallelTreeMaker's constructor (DataParallelTreeMaker.java:167-172). */
    public GpuTreeMaker(ThreadCommSlave comm, GBDTOptimizationParams params, GBDTCoreData trainData,
                        SamplePositionData position, FeatureApprData feaApprData) throws Mp4jException {
        this.comm = comm; this.params = params; this.trainData = trainData; this.position = position;
        this.sampleNum = trainData.sampleNum;
        this.featureDim = trainData.usefulFeatureDim;
        this.nodeOfRow = new int[sampleNum];
        int rank = comm.getRank() * comm.getThreadNum() + comm.getThreadId();
        int world = comm.getSlaveNum() * comm.getThreadNum();
        // rank 0 creates the NCCL id; mp4j carries its 128 bytes to everybody (sum with zeros = copy)
        long[] id = new long[16];
        if (world > 1) {
            if (rank == 0) nativeNcclUniqueId(id);
            id = comm.allreduceArray(id, Operands.LONG_OPERAND(), Operators.Long.SUM, 0, id.length);
        }
        handle = nativeCreate(params.max_depth, params.max_leaf_cnt, params.min_split_samples,
                params.tree_grow_policy.equals(TreeGrowPolicy.LOSSCHG_WISE) ? 1 : 0,
                params.min_child_hessian_sum, params.min_split_loss, params.max_abs_leaf_val,
                params.regularization.l1, params.regularization.l2, params.histogram_pool_capacity,
                params.regularization.learningRate, trainData.numTreeInGroup,
                /*device=*/comm.getThreadId(), rank, world, world > 1 ? id : null);
        nativeSetBins(handle, feaApprData.getFeaSplitValStartIndex(), featureDim, sampleNum);
        int[][] x = trainData.getX();                       // slot indices after FeatureApprData's conversion
        long row0 = 0;
        for (int k = 0; k < trainData.cursor2d; k++) {
            nativeUploadRows(handle, x[k], row0, trainData.realNum[k], trainData.maxFeatureDim);
            row0 += trainData.realNum[k];
        }
    }

    @Override
    public Tree make(int groupId) throws Mp4jException {
        // GBDTOptimizer rewrites gradPairs once per round, before the K make() calls (GBDTOptimizer.java:550-552,598-599)
        if (groupId == 0) {
            long row0 = 0;
            for (int k = 0; k < trainData.cursor2d; k++) {
                nativeSetGrads(handle, trainData.gradPairs[k], row0, trainData.realNum[k]);
                row0 += trainData.realNum[k];
            }
        }
        // initAssistData (DataParallelTreeMaker.java:395-470): row subsample with an unseeded Random,
        // feature subsample from a seed every worker agrees on (allreduce MAX of nanoTime)
        byte[] rowMask = null;
        if (params.subsample < 1.0f) {
            Random rand = new Random();
            rowMask = new byte[sampleNum];
            for (int i = 0; i < sampleNum; i++) rowMask[i] = (byte) (RandomUtils.sampleBinary(rand, params.subsample) ? 1 : 0);
        }
        int[] feaIndex = null;
        if (params.feature_sample_rate < 1.0f) {
            long seed = comm.allreduce(System.nanoTime(), Operands.LONG_OPERAND(), Operators.Long.MAX);
            int[] all = new int[featureDim];
            for (int i = 0; i < featureDim; i++) all[i] = i;
            RandomUtils.shuffle(seed, all);
            int nFeaSam = Math.max(1, Math.round(params.feature_sample_rate * featureDim));
            feaIndex = Arrays.copyOf(all, nFeaSam);
            Arrays.sort(feaIndex);
        }

        int cap = params.max_leaf_cnt > 0 ? 2 * params.max_leaf_cnt - 1 : (1 << (params.max_depth + 1)) - 1;
        int[] parent = new int[cap], left = new int[cap], right = new int[cap];
        int[] fid = new int[cap], lo = new int[cap], hi = new int[cap];
        float[] info = new float[cap], lossChg = new float[cap], hess = new float[cap];
        long[] cnt = new long[cap];
        double[] ts = new double[8];
        int n = nativeMakeTree(handle, groupId, feaIndex, rowMask, parent, left, right, fid, lo, hi,
                               info, lossChg, hess, cnt, ts);

        Tree tree = new Tree();                              // Tree / TreeNode / TreeNodeStat, ids in allocation order
        tree.init();
        for (int nid = 0; nid < n; nid++) {
            if (left[nid] != -1) {
                tree.addChilds(nid);                         // children are the next two ids, as in the native tree
                tree.getNode(nid).setSplit(fid[nid], lo[nid], hi[nid]);
            } else {
                tree.getNode(nid).setLeaf(info[nid]);
            }
            TreeNodeStat st = tree.getNodeStat(nid);
            st.setLossChg(lossChg[nid]); st.setHessSum(hess[nid]); st.setNodeSampleCnt(cnt[nid]);
        }
        for (int i = 0; i < 8; i++) totalMs[i] += ts[i];
        // TreeRefiner (l1) and the score update read the final partition (SamplePositionData)
        if ("l1".equalsIgnoreCase(params.objective) || params.subsample < 1.0f) exportPosition(n, left);
        return tree;
    }

    /** The final partition as SamplePositionData holds it: (node id, row) pairs grouped by node,
     *  rows that were not sampled first with node id -1 (SamplePositionData.java:64-87). */
    private void exportPosition(int nNodes, int[] left) {
        nativeGetPositions(handle, nodeOfRow);
        int[] count = new int[nNodes + 1];                   // slot 0: not sampled
        for (int i = 0; i < sampleNum; i++) count[nodeOfRow[i] + 1]++;
        int[] first = new int[nNodes + 1];
        for (int k = 1; k <= nNodes; k++) first[k] = first[k - 1] + count[k - 1];
        List<Tuple<Integer, Integer>> intervals = position.getNodeSampleInterval();
        intervals.clear();
        for (int nid = 0; nid < nNodes; nid++)
            intervals.add(left[nid] == -1 ? new Tuple<>(first[nid + 1], first[nid + 1] + count[nid + 1]) : new Tuple<>(0, 0));
        int[] pairs = position.getPosAndSampleIndex();
        int[] cursor = Arrays.copyOf(first, first.length);
        for (int i = 0; i < sampleNum; i++) {
            int at = cursor[nodeOfRow[i] + 1]++;
            pairs[2 * at] = nodeOfRow[i];
            pairs[2 * at + 1] = i;
        }
    }

    /** TimeStats.getStats (data/gbdt/TimeStats.java:69-78) prints these; the stage times are CUDA-event
     *  times of the stages on the device and add up to the trees' device time. */
    @Override
    public TimeStats getTotalTimeStats() {
        TimeStats t = new TimeStats();
        t.buildHist = Math.round(totalMs[0]); t.buildDetail[0] = Math.round(totalMs[1]); t.buildDetail[1] = Math.round(totalMs[2]);
        t.initStats = Math.round(totalMs[3] + totalMs[6]);   // row partition: the reference times it under initStats' caller
        t.findBestSplit = Math.round(totalMs[4] + totalMs[5]); t.syncBestSplit = Math.round(totalMs[5]);
        t.totalTime = Math.round(totalMs[7]);
        return t;
    }

    @Override
    public void close() {
        if (handle != 0) { nativeDestroy(handle); handle = 0; }
    }

    // one native per entry point of include/ytk_gbdt.h; failures throw YtkLearnException
    private static native void nativeNcclUniqueId(long[] out16);
    private static native long nativeCreate(int maxDepth, int maxLeafCnt, int minSplitSamples, int policy,
            double minChildHess, double minSplitLoss, double maxAbsLeafVal, double l1, double l2,
            double poolCapacityMB, float lr, int classNum, int device, int rank, int world, long[] ncclId16);
    private static native void nativeSetBins(long h, int[] start, int featureDim, long nRows);
    private static native void nativeUploadRows(long h, int[] xChunk, long row0, long nRows, int stride);
    private static native void nativeSetGrads(long h, float[] gpChunk, long row0, long nRows);
    private static native int nativeMakeTree(long h, int gid, int[] feaIndex, byte[] rowMask,
            int[] parent, int[] left, int[] right, int[] fid, int[] lo, int[] hi,
            float[] info, float[] lossChg, float[] hess, long[] cnt, double[] timeStats8);
    private static native void nativeGetPositions(long h, int[] nodeOfRow);
    private static native void nativeDestroy(long h);
}
