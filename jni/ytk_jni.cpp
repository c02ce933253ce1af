// JNI glue between com.fenbi.ytklearn.optimizer.gbdt.GpuTreeMaker (jni/GpuTreeMaker.java) and the C ABI of
// libytkgbdt.so (include/ytk_gbdt.h): one function per native, no logic of its own.  Arrays are pinned
// with Get/ReleasePrimitiveArrayCritical for the duration of one call (the library retains no host
// pointer); failures become YtkLearnException (exception/YtkLearnException.java), which takes a worker
// down the reference's own exit path (worker/TrainWorker.java:188-196).
//
// Build (needs a JDK; this image has none -- tests/test_jni_syntax.py parses the file against
// jni/stub/jni.h, a stand-in declaring just the JNI names used here):
//   g++ -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -Iinclude jni/ytk_jni.cpp \
//       -Lytk-learn_b200 -lytkgbdt -o libytkgbdt_jni.so
#include <jni.h>

#include <cstring>

#include "ytk_gbdt.h"

namespace {

void throw_ytk(JNIEnv* env) {
    jclass cls = env->FindClass("com/fenbi/ytklearn/exception/YtkLearnException");
    if (cls) env->ThrowNew(cls, ytk_gbdt_last_error());
}

// A primitive array pinned for the scope of one native call; null arrays give null pointers.
template <typename T>
class Pinned {
public:
    Pinned(JNIEnv* env, jarray arr, jint release_mode) : env_(env), arr_(arr), mode_(release_mode) {
        p_ = arr ? static_cast<T*>(env->GetPrimitiveArrayCritical(arr, nullptr)) : nullptr;
    }
    ~Pinned() { if (p_) env_->ReleasePrimitiveArrayCritical(arr_, p_, mode_); }
    T* get() const { return p_; }
    bool failed() const { return arr_ && !p_; }
private:
    JNIEnv* env_; jarray arr_; jint mode_; T* p_;
};

ytk_gbdt* H(jlong h) { return reinterpret_cast<ytk_gbdt*>(static_cast<intptr_t>(h)); }

}  // namespace

#define YTK_JNI(ret, name) extern "C" JNIEXPORT ret JNICALL Java_com_fenbi_ytklearn_optimizer_gbdt_GpuTreeMaker_##name

YTK_JNI(void, nativeNcclUniqueId)(JNIEnv* env, jclass, jlongArray out16) {
    if (!out16 || env->GetArrayLength(out16) < 16) { ytk_gbdt_last_error(); return; }
    Pinned<jlong> p(env, out16, 0);
    if (p.failed()) return;
    if (ytk_gbdt_nccl_unique_id(p.get()) != YTK_OK) throw_ytk(env);
}

YTK_JNI(jlong, nativeCreate)(JNIEnv* env, jclass, jint max_depth, jint max_leaf_cnt, jint min_split_samples, jint policy,
                             jdouble min_child_hess, jdouble min_split_loss, jdouble max_abs_leaf_val, jdouble l1, jdouble l2,
                             jdouble pool_capacity_mb, jfloat lr, jint class_num, jint device, jint rank, jint world,
                             jlongArray nccl_id16) {
    ytk_gbdt_params p;
    std::memset(&p, 0, sizeof(p));
    p.max_depth = max_depth; p.max_leaf_cnt = max_leaf_cnt; p.min_split_samples = min_split_samples; p.grow_policy = policy;
    p.min_child_hessian_sum = min_child_hess; p.min_split_loss = min_split_loss; p.max_abs_leaf_val = max_abs_leaf_val;
    p.l1 = l1; p.l2 = l2; p.histogram_pool_capacity = pool_capacity_mb; p.learning_rate = lr; p.class_num = class_num;
    ytk_gbdt* h = nullptr;
    int rc;
    {
        Pinned<jlong> id(env, nccl_id16, JNI_ABORT);
        if (id.failed()) return 0;
        rc = ytk_gbdt_create(&p, device, rank, world, id.get(), &h);
    }
    if (rc != YTK_OK) { throw_ytk(env); return 0; }
    return static_cast<jlong>(reinterpret_cast<intptr_t>(h));
}

YTK_JNI(void, nativeSetBins)(JNIEnv* env, jclass, jlong h, jintArray start, jint feature_dim, jlong n_rows) {
    int rc;
    {
        Pinned<jint> s(env, start, JNI_ABORT);
        if (s.failed()) return;
        rc = ytk_gbdt_set_bins(H(h), reinterpret_cast<const int32_t*>(s.get()), feature_dim, n_rows);
    }
    if (rc != YTK_OK) throw_ytk(env);
}

YTK_JNI(void, nativeUploadRows)(JNIEnv* env, jclass, jlong h, jintArray x_chunk, jlong row0, jlong n_rows, jint stride) {
    int rc;
    {
        Pinned<jint> x(env, x_chunk, JNI_ABORT);
        if (x.failed()) return;
        rc = ytk_gbdt_upload_rows(H(h), reinterpret_cast<const int32_t*>(x.get()), row0, n_rows, stride);
    }
    if (rc != YTK_OK) throw_ytk(env);
}

YTK_JNI(void, nativeSetGrads)(JNIEnv* env, jclass, jlong h, jfloatArray gp_chunk, jlong row0, jlong n_rows) {
    int rc;
    {
        Pinned<jfloat> g(env, gp_chunk, JNI_ABORT);
        if (g.failed()) return;
        rc = ytk_gbdt_set_grads(H(h), g.get(), row0, n_rows);
    }
    if (rc != YTK_OK) throw_ytk(env);
}

YTK_JNI(jint, nativeMakeTree)(JNIEnv* env, jclass, jlong h, jint gid, jintArray fea_index, jbyteArray row_mask,
                              jintArray parent, jintArray left, jintArray right, jintArray fid, jintArray lo, jintArray hi,
                              jfloatArray info, jfloatArray loss_chg, jfloatArray hess, jlongArray cnt, jdoubleArray ts8) {
    ytk_tree_out t;
    std::memset(&t, 0, sizeof(t));
    t.capacity = env->GetArrayLength(parent);
    const jsize n_fea = fea_index ? env->GetArrayLength(fea_index) : 0;
    ytk_time_stats s;
    std::memset(&s, 0, sizeof(s));
    int rc;
    {
        // make_tree runs for milliseconds and calls nothing in the JVM: holding the arrays critical is fine
        Pinned<jint> fi(env, fea_index, JNI_ABORT);
        Pinned<jbyte> rm(env, row_mask, JNI_ABORT);
        Pinned<jint> a_parent(env, parent, 0), a_left(env, left, 0), a_right(env, right, 0), a_fid(env, fid, 0),
            a_lo(env, lo, 0), a_hi(env, hi, 0);
        Pinned<jfloat> a_info(env, info, 0), a_lc(env, loss_chg, 0), a_hess(env, hess, 0);
        Pinned<jlong> a_cnt(env, cnt, 0);
        if (fi.failed() || rm.failed() || a_parent.failed() || a_left.failed() || a_right.failed() || a_fid.failed() ||
            a_lo.failed() || a_hi.failed() || a_info.failed() || a_lc.failed() || a_hess.failed() || a_cnt.failed())
            return 0;                                        // OutOfMemoryError is already pending
        t.parent = reinterpret_cast<int32_t*>(a_parent.get()); t.left = reinterpret_cast<int32_t*>(a_left.get());
        t.right = reinterpret_cast<int32_t*>(a_right.get()); t.fid = reinterpret_cast<int32_t*>(a_fid.get());
        t.lo = reinterpret_cast<int32_t*>(a_lo.get()); t.hi = reinterpret_cast<int32_t*>(a_hi.get());
        t.node_info = a_info.get(); t.loss_chg = a_lc.get(); t.hess_sum = a_hess.get();
        t.sample_cnt = reinterpret_cast<int64_t*>(a_cnt.get());
        rc = ytk_gbdt_make_tree(H(h), gid, reinterpret_cast<const int32_t*>(fi.get()), n_fea,
                                reinterpret_cast<const uint8_t*>(rm.get()), &t, &s);
    }
    if (rc != YTK_OK) { throw_ytk(env); return 0; }
    if (ts8 && env->GetArrayLength(ts8) >= 8) {
        const jdouble v[8] = {s.build_hist, s.compute_hist, s.communicate_hist, s.init_stats,
                              s.find_best_split, s.sync_best_split, s.partition, s.total};
        env->SetDoubleArrayRegion(ts8, 0, 8, v);
    }
    return t.n_nodes;
}

YTK_JNI(void, nativeGetPositions)(JNIEnv* env, jclass, jlong h, jintArray node_of_row) {
    int rc;
    {
        Pinned<jint> p(env, node_of_row, 0);
        if (p.failed()) return;
        rc = ytk_gbdt_get_positions(H(h), reinterpret_cast<int32_t*>(p.get()));
    }
    if (rc != YTK_OK) throw_ytk(env);
}

YTK_JNI(void, nativeDestroy)(JNIEnv*, jclass, jlong h) { ytk_gbdt_destroy(H(h)); }
