// Haar cascade XML loader: own minimal XML reader (no cv::FileStorage here) for both the old
// "opencv-haar-classifier" files the reference ships (data/haarcascade_frontalface_*.xml, loaded at
// src/file_io.cc:11) and the new "opencv-cascade-classifier" layout OpenCV converts them to.
// Semantics follow SURVEY.md 3.3 / A.4: stage threshold = f32(xml) - 1e-5f, applied at load.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "../../include/ugdet.h"
#include "cascade.h"

namespace ugd {
namespace {

struct Elem {
    std::string name, type_id, text;
    std::vector<std::unique_ptr<Elem>> kids;
    const Elem* child(const char* n) const {
        for (auto& k : kids) if (k->name == n) return k.get();
        return nullptr;
    }
};

struct XmlReader {
    const char* p; const char* end; std::string err;
    bool fail(const char* m) { if (err.empty()) err = m; return false; }
    void skip_ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) ++p; }
    bool starts(const char* s) const { size_t n = strlen(s); return (size_t)(end - p) >= n && !memcmp(p, s, n); }
    bool skip_until(const char* s) {
        size_t n = strlen(s);
        while ((size_t)(end - p) >= n) { if (!memcmp(p, s, n)) { p += n; return true; } ++p; }
        return fail("unterminated comment or declaration");
    }
    // Skips comments / PIs / whitespace between nodes.
    bool skip_misc() {
        for (;;) {
            skip_ws();
            if (starts("<!--")) { if (!skip_until("-->")) return false; }
            else if (starts("<?")) { if (!skip_until("?>")) return false; }
            else if (starts("<!")) { if (!skip_until(">")) return false; }
            else return true;
        }
    }
    bool parse_elem(Elem& e, int depth) {
        if (depth > 64) return fail("XML nested too deeply");
        if (p >= end || *p != '<') return fail("expected '<'");
        ++p;
        const char* s = p;
        while (p < end && *p != ' ' && *p != '>' && *p != '/' && *p != '\n' && *p != '\t' && *p != '\r') ++p;
        e.name.assign(s, p);
        if (e.name.empty()) return fail("empty tag name");
        // attributes
        for (;;) {
            skip_ws();
            if (p >= end) return fail("unterminated tag");
            if (*p == '/') { if (p + 1 < end && p[1] == '>') { p += 2; return true; } return fail("bad '/'"); }
            if (*p == '>') { ++p; break; }
            const char* a = p;
            while (p < end && *p != '=' && *p != ' ' && *p != '>') ++p;
            std::string an(a, p);
            skip_ws();
            if (p >= end || *p != '=') return fail("attribute without value");
            ++p; skip_ws();
            if (p >= end || (*p != '"' && *p != '\'')) return fail("attribute value not quoted");
            char q = *p++;
            const char* v = p;
            while (p < end && *p != q) ++p;
            if (p >= end) return fail("unterminated attribute");
            if (an == "type_id") e.type_id.assign(v, p);
            ++p;
        }
        // content
        for (;;) {
            const char* t = p;
            while (p < end && *p != '<') ++p;
            e.text.append(t, p);
            if (p >= end) return fail("unexpected end of file inside element");
            if (starts("<!--")) { if (!skip_until("-->")) return false; continue; }
            if (starts("</")) {
                p += 2;
                const char* n = p;
                while (p < end && *p != '>') ++p;
                if (p >= end) return fail("unterminated end tag");
                std::string en(n, p);
                while (!en.empty() && (en.back() == ' ' || en.back() == '\n')) en.pop_back();
                ++p;
                if (en != e.name) return fail("mismatched end tag");
                return true;
            }
            std::unique_ptr<Elem> k(new Elem);
            if (!parse_elem(*k, depth + 1)) return false;
            e.kids.push_back(std::move(k));
        }
    }
};

bool parse_numbers(const std::string& s, std::vector<double>& out) {
    out.clear();
    const char* p = s.c_str();
    for (;;) {
        while (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t') ++p;
        if (!*p) return true;
        char* e = nullptr;
        double v = strtod(p, &e);
        if (e == p) return false;
        out.push_back(v);
        p = e;
    }
}

bool one_number(const Elem* e, double& v) {
    if (!e) return false;
    std::vector<double> t;
    if (!parse_numbers(e->text, t) || t.size() != 1) return false;
    v = t[0];
    return true;
}

bool parse_rects(const Elem* rects, HostNode& n, std::string& err) {
    if (!rects) { err = "feature without <rects>"; return false; }
    if (rects->kids.empty() || rects->kids.size() > 3) { err = "feature must have 1..3 rects"; return false; }
    int k = 0;
    for (auto& r : rects->kids) {
        std::vector<double> v;
        if (!parse_numbers(r->text, v) || v.size() != 5) { err = "rect is not 'x y w h weight'"; return false; }
        for (int i = 0; i < 4; ++i) n.rect[k][i] = (int32_t)v[i];
        n.weight[k] = (float)v[4];
        ++k;
    }
    return true;
}

int load_old(const Elem& root, Cascade& c, std::string& err) {
    std::vector<double> v;
    const Elem* sz = root.child("size");
    if (!sz || !parse_numbers(sz->text, v) || v.size() != 2) { err = "missing <size>"; return UGD_ERR_PARSE; }
    c.win_w = (int)v[0]; c.win_h = (int)v[1];
    const Elem* stages = root.child("stages");
    if (!stages) { err = "missing <stages>"; return UGD_ERR_PARSE; }
    for (auto& st : stages->kids) {
        const Elem* trees = st->child("trees");
        double thr;
        if (!trees || !one_number(st->child("stage_threshold"), thr)) { err = "stage without trees/threshold"; return UGD_ERR_PARSE; }
        c.stage_first_tree.push_back(c.ntrees());
        for (auto& tr : trees->kids) {
            c.tree_first_node.push_back(c.nnodes());
            c.tree_nnodes.push_back((int32_t)tr->kids.size());
            for (auto& nd : tr->kids) {
                HostNode n{};
                const Elem* feat = nd->child("feature");
                if (!feat) { err = "node without <feature>"; return UGD_ERR_PARSE; }
                if (!parse_rects(feat->child("rects"), n, err)) return UGD_ERR_PARSE;
                double t = 0;
                if (one_number(feat->child("tilted"), t) && t != 0) n.tilted = 1;
                double d;
                if (!one_number(nd->child("threshold"), d)) { err = "node without <threshold>"; return UGD_ERR_PARSE; }
                n.thr = (float)d;
                if (one_number(nd->child("left_val"), d)) n.left_val = (float)d;
                else if (one_number(nd->child("left_node"), d)) n.left = (int32_t)d;
                else { err = "node without left_val/left_node"; return UGD_ERR_PARSE; }
                if (one_number(nd->child("right_val"), d)) n.right_val = (float)d;
                else if (one_number(nd->child("right_node"), d)) n.right = (int32_t)d;
                else { err = "node without right_val/right_node"; return UGD_ERR_PARSE; }
                c.nodes.push_back(n);
            }
        }
        c.stage_ntrees.push_back(c.ntrees() - c.stage_first_tree.back());
        c.stage_thr.push_back((float)thr - 1e-5f);
        // the stage tree (only the OpenCV-2.4 legacy mode walks it; 4.x evaluates the stages in file order)
        double par = (double)c.nstages() - 2.0, nxt = -1.0;
        one_number(st->child("parent"), par);
        one_number(st->child("next"), nxt);
        c.stage_thr_raw.push_back((float)thr);
        c.stage_parent.push_back((int32_t)par);
        c.stage_next.push_back((int32_t)nxt);
    }
    for (int s = 0; s < c.nstages(); ++s)
        if (c.stage_parent[s] >= s || c.stage_parent[s] < -1 || c.stage_next[s] >= c.nstages() || c.stage_next[s] < -1 ||
            (c.stage_next[s] >= 0 && c.stage_next[s] <= s)) { err = "malformed stage tree (<parent>/<next>)"; return UGD_ERR_PARSE; }
    c.old_format = true;
    return UGD_OK;
}

int load_new(const Elem& root, Cascade& c, std::string& err) {
    const Elem* ft = root.child("featureType");
    if (!ft || ft->text.find("HAAR") == std::string::npos) { err = "only HAAR feature cascades are supported"; return UGD_ERR_UNSUPPORTED; }
    double w, h;
    if (!one_number(root.child("width"), w) || !one_number(root.child("height"), h)) { err = "missing width/height"; return UGD_ERR_PARSE; }
    c.win_w = (int)w; c.win_h = (int)h;
    const Elem* feats = root.child("features");
    const Elem* stages = root.child("stages");
    if (!feats || !stages) { err = "missing <features>/<stages>"; return UGD_ERR_PARSE; }
    std::vector<HostNode> fl;
    for (auto& f : feats->kids) {
        HostNode n{};
        if (!parse_rects(f->child("rects"), n, err)) return UGD_ERR_PARSE;
        double t = 0;
        if (one_number(f->child("tilted"), t) && t != 0) n.tilted = 1;
        fl.push_back(n);
    }
    std::vector<double> in, lv;
    for (auto& st : stages->kids) {
        double thr;
        const Elem* wcs = st->child("weakClassifiers");
        if (!wcs || !one_number(st->child("stageThreshold"), thr)) { err = "stage without weakClassifiers/threshold"; return UGD_ERR_PARSE; }
        c.stage_first_tree.push_back(c.ntrees());
        for (auto& wc : wcs->kids) {
            const Elem* a = wc->child("internalNodes");
            const Elem* b = wc->child("leafValues");
            if (!a || !b || !parse_numbers(a->text, in) || !parse_numbers(b->text, lv) || in.empty() || in.size() % 4) {
                err = "bad weak classifier"; return UGD_ERR_PARSE;
            }
            int nn = (int)in.size() / 4;
            c.tree_first_node.push_back(c.nnodes());
            c.tree_nnodes.push_back(nn);
            for (int k = 0; k < nn; ++k) {
                int l = (int)in[4 * k], r = (int)in[4 * k + 1], fi = (int)in[4 * k + 2];
                if (fi < 0 || fi >= (int)fl.size() || (l <= 0 && -l >= (int)lv.size()) || (r <= 0 && -r >= (int)lv.size())) {
                    err = "weak classifier index out of range"; return UGD_ERR_PARSE;
                }
                HostNode n = fl[fi];
                n.thr = (float)in[4 * k + 3];
                n.left = l > 0 ? l : 0; n.right = r > 0 ? r : 0;
                n.left_val = l <= 0 ? (float)lv[-l] : 0.f;
                n.right_val = r <= 0 ? (float)lv[-r] : 0.f;
                c.nodes.push_back(n);
            }
        }
        c.stage_ntrees.push_back(c.ntrees() - c.stage_first_tree.back());
        c.stage_thr.push_back((float)thr - 1e-5f);
        c.stage_thr_raw.push_back((float)thr);
        c.stage_parent.push_back(c.nstages() - 2);        // the new format has no stage tree: a linear chain
        c.stage_next.push_back(-1);
    }
    c.old_format = false;
    return UGD_OK;
}

// lowest set bit exponent of a finite non-zero float (value = m * 2^e with m odd -> e)
int lsb_exponent(float f) {
    int e; float m = frexpf(fabsf(f), &e);          // f = m * 2^e, m in [0.5,1)
    uint32_t mant = (uint32_t)ldexpf(m, 24);        // 24-bit integer
    int tz = 0; while (!(mant & 1u)) { mant >>= 1; ++tz; }
    return e - 24 + tz;
}

int validate_and_derive(Cascade& c, std::string& err) {
    if (c.win_w < 4 || c.win_h < 4 || c.win_w > 64 || c.win_h > 64) { err = "window size outside 4..64"; return UGD_ERR_UNSUPPORTED; }
    if (c.nstages() == 0 || c.nstages() > 32000) { err = "cascade has no stages"; return UGD_ERR_PARSE; }
    c.max_tree_nodes = 0;
    int min_lsb = 1000; double max_bound = 0;
    c.stage_first_node.assign(c.nstages() + 1, 0);
    c.corners.resize(c.nodes.size());
    for (int s = 0; s < c.nstages(); ++s) {
        double bound = 0;
        if (c.stage_ntrees[s] <= 0) { err = "stage without weak classifiers"; return UGD_ERR_PARSE; }
        c.stage_first_node[s] = c.tree_first_node[c.stage_first_tree[s]];
        for (int t = c.stage_first_tree[s]; t < c.stage_first_tree[s] + c.stage_ntrees[s]; ++t) {
            int nn = c.tree_nnodes[t], base = c.tree_first_node[t];
            if (nn <= 0 || nn > 255) { err = "tree with 0 or more than 255 nodes"; return UGD_ERR_UNSUPPORTED; }
            if (nn > c.max_tree_nodes) c.max_tree_nodes = nn;
            double tmax = 0;
            for (int k = 0; k < nn; ++k) {
                HostNode& n = c.nodes[base + k];
                if (n.left < 0 || n.right < 0 || n.left >= nn || n.right >= nn || (n.left && n.left <= k) || (n.right && n.right <= k)) {
                    err = "tree child index must follow its parent"; return UGD_ERR_UNSUPPORTED;
                }
                int sumw = 0;
                for (int r = 0; r < 3; ++r) {
                    float w = n.weight[r];
                    if (w != (float)(int)w || fabsf(w) > 127.f) { err = "non-integer feature weight"; return UGD_ERR_UNSUPPORTED; }
                    if (w == 0.f) continue;
                    const int32_t* q = n.rect[r];
                    const bool inside = n.tilted
                        ? (q[0] - q[3] >= 0 && q[1] >= 0 && q[2] > 0 && q[3] > 0 && q[0] + q[2] <= c.win_w && q[1] + q[2] + q[3] <= c.win_h)
                        : (q[0] >= 0 && q[1] >= 0 && q[2] > 0 && q[3] > 0 && q[0] + q[2] <= c.win_w && q[1] + q[3] <= c.win_h);
                    if (!inside) { err = "feature rectangle outside the window"; return UGD_ERR_PARSE; }
                    sumw += abs((int)w);
                }
                if (n.weight[0] == 0.f || n.weight[1] == 0.f) { err = "feature needs at least two rects"; return UGD_ERR_UNSUPPORTED; }
                // int32 evaluation is exact iff every partial sum stays below 2^24 (float mantissa)
                if ((double)sumw * c.win_w * c.win_h * 255.0 >= 16777216.0) { err = "feature weights too large for exact evaluation"; return UGD_ERR_UNSUPPORTED; }
                {   // corner list: merge duplicate corners, drop zero coefficients
                    CornerNode cn{};
                    int m = 0;
                    int tx[12], ty[12], tc[12];
                    for (int r = 0; r < 3; ++r) {
                        int w = (int)n.weight[r];
                        if (!w) continue;
                        const int32_t* q = n.rect[r];
                        // upright: S(x,y) - S(x+w,y) - S(x,y+h) + S(x+w,y+h)
                        // tilted : T(x,y) - T(x-h,y+h) - T(x+w,y+w) + T(x+w-h,y+w+h)   (cv2's CV_TILTED_OFS)
                        const int ux[4] = {q[0], q[0] + q[2], q[0], q[0] + q[2]};
                        const int uy[4] = {q[1], q[1], q[1] + q[3], q[1] + q[3]};
                        const int tx4[4] = {q[0], q[0] - q[3], q[0] + q[2], q[0] + q[2] - q[3]};
                        const int ty4[4] = {q[1], q[1] + q[3], q[1] + q[2], q[1] + q[2] + q[3]};
                        const int* px = n.tilted ? tx4 : ux;
                        const int* py = n.tilted ? ty4 : uy;
                        const int sg[4] = {w, -w, -w, w};
                        for (int e = 0; e < 4; ++e) {
                            int f = 0;
                            for (; f < m; ++f) if (tx[f] == px[e] && ty[f] == py[e]) break;
                            if (f == m) { tx[m] = px[e]; ty[m] = py[e]; tc[m] = 0; ++m; }
                            tc[f] += sg[e];
                        }
                    }
                    int o = 0;
                    for (int f = 0; f < m; ++f) {
                        if (!tc[f]) continue;
                        cn.cx[o] = (int8_t)tx[f]; cn.cy[o] = (int8_t)ty[f]; cn.coef[o] = (int8_t)tc[f]; ++o;
                    }
                    cn.nc = o <= 6 ? 6 : (o <= 8 ? 8 : (o <= 9 ? 9 : 12));   // padding entries: corner (0,0), coef 0
                    cn.plane = n.tilted ? 1 : 0;
                    if (n.tilted) c.has_tilted = true;
                    c.corners[base + k] = cn;
                }
                for (int side = 0; side < 2; ++side) {
                    if ((side ? n.right : n.left) != 0) continue;
                    float lf = side ? n.right_val : n.left_val;
                    if (!std::isfinite(lf)) { err = "non-finite leaf value"; return UGD_ERR_PARSE; }
                    if (lf != 0.f) { int e = lsb_exponent(lf); if (e < min_lsb) min_lsb = e; }
                    if (fabs((double)lf) > tmax) tmax = fabs((double)lf);
                }
            }
            bound += tmax;
        }
        if (bound > max_bound) max_bound = bound;
    }
    c.stage_first_node[c.nstages()] = c.nnodes();
    // all leaves are multiples of 2^min_lsb and |stage sum| <= max_bound: exact in fp64 iff it fits 53 bits
    int need = (max_bound > 0 ? (int)ceil(log2(max_bound + 1)) : 0) - (min_lsb == 1000 ? 0 : min_lsb);
    c.sum_exact = need <= 52;
    // the stump evaluator sums base = sum(right) plus (left - right) of the nodes that go left: every
    // partial sum is bounded by 3 * max_bound, still a multiple of 2^min_lsb
    int need3 = (max_bound > 0 ? (int)ceil(log2(3.0 * max_bound + 1)) : 0) - (min_lsb == 1000 ? 0 : min_lsb);
    c.delta_exact = c.sum_exact && need3 <= 52;
    return UGD_OK;
}

}  // namespace

int load_cascade_xml(const std::string& path, Cascade& out, std::string& err) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) { err = "cannot open " + path; return UGD_ERR_IO; }
    std::string buf;
    char tmp[1 << 16];
    size_t n;
    while ((n = fread(tmp, 1, sizeof tmp, f)) > 0) buf.append(tmp, n);
    fclose(f);
    XmlReader xr{buf.data(), buf.data() + buf.size(), {}};
    if (!xr.skip_misc()) { err = xr.err; return UGD_ERR_PARSE; }
    Elem root;
    if (!xr.parse_elem(root, 0)) { err = "XML: " + xr.err; return UGD_ERR_PARSE; }
    if (root.name != "opencv_storage") { err = "not an opencv_storage file"; return UGD_ERR_PARSE; }
    int rc = UGD_ERR_PARSE;
    err = "no Haar cascade node in file";
    for (auto& k : root.kids) {
        if (k->type_id == "opencv-haar-classifier") { err.clear(); rc = load_old(*k, out, err); break; }
        // some shipped files (haarcascade_frontalcatface*.xml) carry no type_id on <cascade>
        if (k->type_id == "opencv-cascade-classifier" || (k->type_id.empty() && k->child("stageType"))) { err.clear(); rc = load_new(*k, out, err); break; }
    }
    if (rc != UGD_OK) return rc;
    return validate_and_derive(out, err);
}

}  // namespace ugd
