// the text formats either side of the tree maker, host only (no device code here):
//   * the model text GBDTModel.dumpModel / loadModel write and read
//     (J/data/gbdt/GBDTModel.java:63-97, J/data/gbdt/Tree.java:192-291, patterns :47-48, TreeNodeStat.print
//     J/data/gbdt/TreeNodeStat.java:71-79) -- byte for byte, including Java's Float.toString;
//   * the ytklearn training-row text format (docs/data_format.md:1-40; CoreData.line2FeatureMap
//     J/dataflow/CoreData.java:369-378, GBDTCoreData.trainDataSplit / yExtract
//     J/dataflow/GBDTCoreData.java:205-249) parsed straight into the CSR arrays ytk_gbdt_quantize_csr takes.
// J/ = /root/reference/src/main/java/com/fenbi/ytklearn/.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "ytk_gbdt.h"

namespace ytk { void set_last_error(const char* msg); }

namespace {

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    ytk::set_last_error(buf);
    return code;
}

// java.lang.Float.toString: the shortest decimal that reads back as the same float -- at least two digits,
// the closest such decimal (so Float.MIN_VALUE is "1.4E-45") -- laid out as "ddd.ddd" for 1e-3 <= |v| < 1e7 and
// as computerised scientific notation "d.dddE[-]n" otherwise.  This is the rule the JDK documents since 19;
// JDK 8's FloatingDecimal prints one digit more for a few values (JDK-4511638, e.g. Float.MIN_NORMAL), which
// reads back as the same float.
std::string java_float(float v) {
    if (v != v) return "NaN";
    if (std::isinf(v)) return v > 0 ? "Infinity" : "-Infinity";
    if (v == 0.0f) return std::signbit(v) ? "-0.0" : "0.0";
    char buf[48];
    int prec = 2;
    for (; prec <= 9; prec++) {
        snprintf(buf, sizeof(buf), "%.*e", prec - 1, (double)v);
        if (strtof(buf, nullptr) == v) break;
    }
    // buf = [-]d[.ddd]e[+-]xx
    std::string s(buf);
    const bool neg = s[0] == '-';
    if (neg) s.erase(0, 1);
    const size_t epos = s.find('e');
    const int exp10 = atoi(s.c_str() + epos + 1);
    std::string digits;
    for (size_t i = 0; i < epos; i++) if (s[i] != '.') digits.push_back(s[i]);
    while (digits.size() > 1 && digits.back() == '0') digits.pop_back();
    std::string out = neg ? "-" : "";
    const double a = std::fabs((double)v);
    if (a >= 1e-3 && a < 1e7) {
        if (exp10 >= 0) {
            std::string ip = digits.substr(0, std::min<size_t>(digits.size(), (size_t)exp10 + 1));
            while ((int)ip.size() < exp10 + 1) ip.push_back('0');
            std::string fp = (int)digits.size() > exp10 + 1 ? digits.substr(exp10 + 1) : "0";
            out += ip + "." + fp;
        } else {
            out += "0." + std::string((size_t)(-exp10 - 1), '0') + digits;
        }
    } else {
        out += digits.substr(0, 1) + "." + (digits.size() > 1 ? digits.substr(1) : "0") + "E" + std::to_string(exp10);
    }
    return out;
}

struct Writer {
    char* out; int64_t cap; int64_t n = 0;
    void put(const std::string& s) {
        for (char c : s) { if (out && n < cap) out[n] = c; n++; }
    }
    int64_t finish() { if (out && cap > 0) out[std::min(n, cap - 1)] = 0; return n; }
};

int max_depth_to_leaf(const ytk_tree_out* t, int nid) {              // Tree.maxDepthToLeaf (:84-90)
    if (t->left[nid] == -1) return 0;
    return std::max(max_depth_to_leaf(t, t->left[nid]), max_depth_to_leaf(t, t->right[nid])) + 1;
}

void dump_node(const ytk_tree_out* t, const float* thr, const int32_t* dflt, const char* const* names, int nid, int depth,
               bool stats, Writer& w) {                               // Tree.dump (:269-291)
    w.put(std::string((size_t)depth, '\t'));
    char num[64];
    if (t->left[nid] == -1) {
        w.put(std::to_string(nid) + ":leaf=" + java_float(t->node_info[nid]));
        if (stats) {
            snprintf(num, sizeof(num), "%lld", (long long)t->sample_cnt[nid]);
            w.put(",hess_sum=" + java_float(t->hess_sum[nid]) + ",sample_cnt=" + num);
        }
        w.put("\n");
        return;
    }
    const std::string name = names ? std::string(names[t->fid[nid]]) : std::to_string(t->fid[nid]);
    const int missing = dflt ? dflt[nid] : t->left[nid];              // TreeNode.isDefaultLeft starts true (:56)
    w.put(std::to_string(nid) + ":[f_" + name + "<=" + java_float(thr[nid]) + "] yes=" + std::to_string(t->left[nid]) + ",no=" +
          std::to_string(t->right[nid]) + ",missing=" + std::to_string(missing));
    if (stats) {
        snprintf(num, sizeof(num), "%lld", (long long)t->sample_cnt[nid]);
        w.put(",gain=" + java_float(t->loss_chg[nid]) + ",hess_sum=" + java_float(t->hess_sum[nid]) + ",sample_cnt=" + num);
    }
    w.put("\n");
    dump_node(t, thr, dflt, names, t->left[nid], depth + 1, stats, w);
    dump_node(t, thr, dflt, names, t->right[nid], depth + 1, stats, w);
}

// one line of `text` starting at *pos (without its terminator); false at the end
bool next_line(const char* text, int64_t len, int64_t* pos, std::string* line) {
    if (*pos >= len) return false;
    int64_t e = *pos;
    while (e < len && text[e] != '\n') e++;
    line->assign(text + *pos, (size_t)(e - *pos));
    if (!line->empty() && line->back() == '\r') line->pop_back();
    *pos = e < len ? e + 1 : e;
    return true;
}

std::string after(const std::string& s, const char* key) {           // the value after "key", up to ',' ' ' or ']'
    const size_t p = s.find(key);
    if (p == std::string::npos) return "";
    size_t b = p + strlen(key), e = b;
    while (e < s.size() && s[e] != ',' && s[e] != ' ' && s[e] != ']' && s[e] != '\t') e++;
    return s.substr(b, e - b);
}

std::string trim(const std::string& s) {                              // String.trim: code points <= ' '
    size_t b = 0, e = s.size();
    while (b < e && (unsigned char)s[b] <= ' ') b++;
    while (e > b && (unsigned char)s[e - 1] <= ' ') e--;
    return s.substr(b, e - b);
}

// String.split(delim) with a literal delimiter: trailing empty strings are dropped as Java does
std::vector<std::string> split(const std::string& s, const std::string& d) {
    std::vector<std::string> out;
    size_t b = 0;
    for (;;) {
        const size_t p = d.empty() ? std::string::npos : s.find(d, b);
        if (p == std::string::npos) { out.push_back(s.substr(b)); break; }
        out.push_back(s.substr(b, p - b));
        b = p + d.size();
    }
    while (!out.empty() && out.back().empty()) out.pop_back();
    return out;
}

bool parse_float(const std::string& s, float* v) {                    // Float.parseFloat (leading/trailing blanks allowed)
    const std::string t = trim(s);
    if (t.empty()) return false;
    char* end = nullptr;
    *v = strtof(t.c_str(), &end);
    if (end == t.c_str()) return false;
    while (*end == 'f' || *end == 'F' || *end == 'd' || *end == 'D') end++;
    return *end == 0;
}

}  // namespace

extern "C" {

int ytk_gbdt_float_to_string(float v, char* out, int32_t cap) {
    if (!out || cap < 1) return fail(YTK_ERR_ARG, "float_to_string: bad buffer");
    const std::string s = java_float(v);
    if ((int)s.size() + 1 > cap) return fail(YTK_ERR_CAPACITY, "float_to_string: needs %d bytes", (int)s.size() + 1);
    memcpy(out, s.c_str(), s.size() + 1);
    return YTK_OK;
}

int64_t ytk_gbdt_dump_model_header(float base_prediction, int32_t class_num, const char* loss_function, int32_t tree_num,
                                   char* out, int64_t cap) {
    Writer w{out, cap};
    // "uniform_base_prediction=" + basePrediction: string concatenation of a float = Float.toString
    w.put("uniform_base_prediction=" + java_float(base_prediction) + "\n");
    w.put("class_num=" + std::to_string(class_num) + "\n");
    w.put(std::string("loss_function=") + (loss_function ? loss_function : "") + "\n");
    w.put("tree_num=" + std::to_string(tree_num) + "\n");
    return w.finish();
}

int64_t ytk_gbdt_dump_tree(const ytk_tree_out* tree, const float* threshold, const int32_t* default_child,
                           const char* const* feature_names, int32_t iter, int32_t with_stats, char* out, int64_t cap) {
    if (!tree || !threshold || tree->n_nodes < 1) { fail(YTK_ERR_ARG, "[GBDT] save invalid model, nodeNum=%d should > 0", tree ? tree->n_nodes : 0); return -1; }
    for (int i = 0; i < tree->n_nodes; i++) {
        if (tree->left[i] == -1) continue;
        if (tree->left[i] < 0 || tree->left[i] >= tree->n_nodes || tree->right[i] < 0 || tree->right[i] >= tree->n_nodes) {
            fail(YTK_ERR_ARG, "dump_tree: node %d has children outside the tree", i);
            return -1;
        }
    }
    Writer w{out, cap};
    int leaves = 0;
    for (int i = 0; i < tree->n_nodes; i++) leaves += tree->left[i] == -1;      // every allocated node is reachable
    char head[128];
    snprintf(head, sizeof(head), "booster[%d] depth=%d,node_num=%d,leaf_cnt=%d\n", iter + 1, max_depth_to_leaf(tree, 0),
             tree->n_nodes, leaves);
    w.put(head);
    dump_node(tree, threshold, default_child, feature_names, 0, 0, with_stats != 0, w);
    return w.finish();
}

int ytk_gbdt_load_model_header(const char* text, int64_t len, float* base_prediction, int32_t* class_num, char* loss_function,
                               int32_t loss_cap, int32_t* tree_num, int64_t* consumed) {
    if (!text || !base_prediction || !class_num || !tree_num || !consumed) return fail(YTK_ERR_ARG, "load_model_header: null argument");
    int64_t pos = 0;
    std::string line[4];
    for (int i = 0; i < 4; i++) {
        if (!next_line(text, len, &pos, &line[i])) return fail(YTK_ERR_ARG, "GBDT: load model error, header line %d missing", i + 1);
        const size_t eq = line[i].find('=');
        if (eq == std::string::npos) return fail(YTK_ERR_ARG, "GBDT: load model error, header line: %s", line[i].c_str());
        line[i] = line[i].substr(eq + 1);               // split("=")[1]
    }
    if (!parse_float(line[0], base_prediction)) return fail(YTK_ERR_ARG, "GBDT: load model error, base prediction: %s", line[0].c_str());
    *class_num = atoi(line[1].c_str());
    if (loss_function && loss_cap > 0) snprintf(loss_function, (size_t)loss_cap, "%s", line[2].c_str());
    *tree_num = atoi(line[3].c_str());
    if (*tree_num == 0) return fail(YTK_ERR_ARG, "GBDT: load model error, tree number is 0!");
    *consumed = pos;
    return YTK_OK;
}

int ytk_gbdt_load_tree(const char* text, int64_t len, ytk_tree_out* tree, float* threshold, int32_t* default_child, char* names,
                       int32_t name_stride, int64_t* consumed) {
    if (!text || !tree || !threshold || !consumed) return fail(YTK_ERR_ARG, "load_tree: null argument");
    int64_t pos = 0;
    std::string line;
    if (!next_line(text, len, &pos, &line)) return fail(YTK_ERR_ARG, "[GBDT] load invalid model: no booster line");
    // "booster[i] depth=d,node_num=n,leaf_cnt=l": split(",")[1].split("=")[1]
    const std::vector<std::string> head = split(line, ",");
    if (head.size() < 2 || head[1].find('=') == std::string::npos) return fail(YTK_ERR_ARG, "[GBDT] load invalid model, line: %s", line.c_str());
    const int n = atoi(head[1].substr(head[1].find('=') + 1).c_str());
    if (n <= 0) return fail(YTK_ERR_ARG, "[GBDT] load invalid model, nodeNum=%d should > 0", n);
    if (n > tree->capacity) return fail(YTK_ERR_CAPACITY, "load_tree: %d nodes, capacity %d", n, tree->capacity);
    tree->n_nodes = n;
    for (int i = 0; i < n; i++) {
        tree->parent[i] = -1; tree->left[i] = -1; tree->right[i] = -1; tree->fid[i] = -1; tree->lo[i] = 0; tree->hi[i] = 0;
        tree->node_info[i] = 0.f; tree->loss_chg[i] = 0.f; tree->hess_sum[i] = 0.f; tree->sample_cnt[i] = 0;
        threshold[i] = 0.f;
        if (default_child) default_child[i] = -1;
        if (names && name_stride > 0) names[(size_t)i * name_stride] = 0;
    }
    for (int k = 0; k < n; k++) {
        if (!next_line(text, len, &pos, &line)) return fail(YTK_ERR_ARG, "[GBDT] parse model error, %d of %d node lines", k, n);
        const std::string s = trim(line);
        const size_t colon = s.find(':');
        if (colon == std::string::npos) return fail(YTK_ERR_ARG, "[GBDT] parse model error, line:%s", line.c_str());
        const int nid = atoi(s.substr(0, colon).c_str());
        if (nid < 0 || nid >= n) return fail(YTK_ERR_ARG, "[GBDT] parse model error, node id in line:%s", line.c_str());
        if (s.find("leaf") != std::string::npos) {                       // parseLeaf (:218-229)
            float v = 0.f, hs = 0.f;
            if (!parse_float(after(s, "leaf="), &v)) return fail(YTK_ERR_ARG, "[GBDT] parse model error, leaf line:%s", line.c_str());
            tree->node_info[nid] = v;
            if (parse_float(after(s, "hess_sum="), &hs)) tree->hess_sum[nid] = hs;
            tree->sample_cnt[nid] = atoll(after(s, "sample_cnt=").c_str());
        } else {                                                          // parseInnerNode (:231-256)
            const size_t fb = s.find("[f_"), le = s.find("<=", fb == std::string::npos ? 0 : fb);
            if (fb == std::string::npos || le == std::string::npos) return fail(YTK_ERR_ARG, "[GBDT] parse model error, non-leaf line:%s", line.c_str());
            const std::string name = s.substr(fb + 3, le - fb - 3);
            float cond = 0.f, gain = 0.f, hs = 0.f;
            if (!parse_float(after(s.substr(le), "<="), &cond)) return fail(YTK_ERR_ARG, "[GBDT] parse model error, non-leaf line:%s", line.c_str());
            const int l = atoi(after(s, "yes=").c_str()), r = atoi(after(s, "no=").c_str()), miss = atoi(after(s, "missing=").c_str());
            if (l < 0 || l >= n || r < 0 || r >= n) return fail(YTK_ERR_ARG, "[GBDT] parse model error, children in line:%s", line.c_str());
            tree->left[nid] = l; tree->right[nid] = r;
            tree->parent[l] = nid; tree->parent[r] = nid;
            threshold[nid] = cond;
            if (default_child) default_child[nid] = miss == l ? l : r;   // isDefaultLeft = (missing == leftChild)
            char* endp = nullptr;
            const long f = strtol(name.c_str(), &endp, 10);
            tree->fid[nid] = (endp != name.c_str() && *endp == 0) ? (int32_t)f : -1;      // a numeric name is its own index
            if (names && name_stride > 0) snprintf(names + (size_t)nid * name_stride, (size_t)name_stride, "%s", name.c_str());
            if (parse_float(after(s, "gain="), &gain)) tree->loss_chg[nid] = gain;
            if (parse_float(after(s, "hess_sum="), &hs)) tree->hess_sum[nid] = hs;
            tree->sample_cnt[nid] = atoll(after(s, "sample_cnt=").c_str());
        }
    }
    *consumed = pos;
    return YTK_OK;
}

int ytk_gbdt_parse_rows(const char* text, int64_t len, const char* x_delim, const char* y_delim, const char* features_delim,
                        const char* name_val_delim, int32_t class_num, const char* const* dict_sorted, int32_t n_dict,
                        int64_t max_rows, int64_t max_entries, float* weight, float* y, float* init_score, int64_t* row_ptr,
                        int32_t* fids, int32_t* value_bits, int64_t* n_rows, int64_t* n_entries, int64_t* consumed) {
    if (!text || !x_delim || !y_delim || !features_delim || !name_val_delim || !weight || !y || !row_ptr || !n_rows || !n_entries ||
        !consumed || class_num < 1 || (max_entries > 0 && (!fids || !value_bits)))
        return fail(YTK_ERR_ARG, "parse_rows: bad arguments");
    const std::string xd(x_delim), yd(y_delim), fd(features_delim), nd(name_val_delim);
    int64_t pos = 0, rows = 0, ents = 0;
    std::string line;
    std::vector<std::pair<int32_t, int32_t>> row;
    row_ptr[0] = 0;
    while (rows < max_rows) {
        const int64_t line_start = pos;
        if (!next_line(text, len, &pos, &line)) break;
        if (pos == len && len > 0 && text[len - 1] != '\n') { pos = line_start; break; }   // an unterminated last line: the caller's next chunk
        const std::string tl = trim(line);
        if (tl.empty()) continue;
        const std::vector<std::string> info = split(tl, xd);                     // trainDataSplit (:205-209)
        if (info.size() < 3) return fail(YTK_ERR_ARG, "[GBDT] data format error! line:%s", line.c_str());
        float wv;
        if (!parse_float(info[0], &wv)) return fail(YTK_ERR_ARG, "[GBDT] data format error (weight)! line:%s", line.c_str());
        float* yr = y + rows * class_num;
        if (class_num == 1) {
            if (!parse_float(info[1], &yr[0])) return fail(YTK_ERR_ARG, "[GBDT] label error, line: %s", line.c_str());
        } else {                                                                  // yExtract, softmax branch (:227-249)
            const std::vector<std::string> li = split(info[1], yd);
            if ((int)li.size() != class_num && li.size() != 1)
                return fail(YTK_ERR_ARG, "[GBDT] label num must equal %d or 1, line: %s", class_num, line.c_str());
            if (li.size() == 1) {
                const int clazz = atoi(trim(li[0]).c_str());
                if (clazz < 0 || clazz >= class_num) return fail(YTK_ERR_ARG, "multi classification label must in range [0,K-1]!\n%s", line.c_str());
                for (int k = 0; k < class_num; k++) yr[k] = 0.f;
                yr[clazz] = 1.0f;
            } else {
                for (int k = 0; k < class_num; k++)
                    if (!parse_float(li[k], &yr[k])) return fail(YTK_ERR_ARG, "[GBDT] label error, line: %s", line.c_str());
            }
        }
        row.clear();
        for (const std::string& f : split(trim(info[2]), fd)) {                  // line2FeatureMap (:369-378)
            const std::vector<std::string> fv = split(f, nd);
            if (fv.size() < 2) return fail(YTK_ERR_ARG, "[GBDT] data format error! feature '%s' in line:%s", f.c_str(), line.c_str());
            float v;
            if (!parse_float(fv[1], &v) || std::isinf(v) || v != v)
                return fail(YTK_ERR_ARG, "data format error, feature value should not be -Infinity, Infinity or NaN; line:%s", line.c_str());
            const std::string name = trim(fv[0]);
            int32_t idx = -1;
            if (dict_sorted) {                                                    // featureMap.getIndex(fn); unknown names are skipped
                int lo = 0, hi = n_dict - 1;
                while (lo <= hi) {
                    const int mid = (lo + hi) / 2, c = strcmp(dict_sorted[mid], name.c_str());
                    if (c == 0) { idx = mid; break; }
                    if (c < 0) lo = mid + 1; else hi = mid - 1;
                }
            } else {                                                              // no dictionary: the name is the index
                char* endp = nullptr;
                const long q = strtol(name.c_str(), &endp, 10);
                if (endp != name.c_str() && *endp == 0 && q >= 0) idx = (int32_t)q;
            }
            if (idx < 0) continue;
            int32_t bits;
            memcpy(&bits, &v, 4);                                                 // Float.floatToRawIntBits(fv) (:381)
            row.push_back({idx, bits});
        }
        // a HashMap keeps the LAST value of a repeated name; rows go out sorted by feature
        std::stable_sort(row.begin(), row.end(), [](const std::pair<int32_t, int32_t>& a, const std::pair<int32_t, int32_t>& b) { return a.first < b.first; });
        size_t kept = 0;
        for (size_t i = 0; i < row.size(); i++) {
            if (i + 1 < row.size() && row[i + 1].first == row[i].first) continue;
            row[kept++] = row[i];
        }
        if (ents + (int64_t)kept > max_entries) { pos = line_start; break; }     // the caller's buffers are full: stop before this row
        for (size_t i = 0; i < kept; i++) { fids[ents] = row[i].first; value_bits[ents] = row[i].second; ents++; }
        if (init_score) {
            for (int k = 0; k < class_num; k++) init_score[rows * class_num + k] = 0.f;
            if (info.size() >= 4) {
                const std::vector<std::string> ii = split(info[3], yd);
                for (int k = 0; k < class_num && k < (int)ii.size(); k++)
                    if (!parse_float(ii[k], &init_score[rows * class_num + k]))
                        return fail(YTK_ERR_ARG, "[GBDT] data format error (initial prediction)! line:%s", line.c_str());
            }
        }
        weight[rows] = wv;
        rows++;
        row_ptr[rows] = ents;
    }
    *n_rows = rows; *n_entries = ents; *consumed = pos;
    return YTK_OK;
}

}  // extern "C"
