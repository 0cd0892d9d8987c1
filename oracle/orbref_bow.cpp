// ORACLE -- TEST INFRASTRUCTURE ONLY.  Never linked into or called by the product path.
//
// CPU restatement of the reference's bag-of-words transform (DBoW2, vendored under
// ORB_SLAM2/Thirdparty/DBoW2/DBoW2): the vocabulary text loader, the tree descent and the BowVector /
// FeatureVector assembly, on flat arrays with the PODs of include/orbgpu.h.  Pinned against the reference's own
// DBoW2 compiled unmodified (oracle/_ref/libdbow2_ref.so, tests/test_bow_cpu.py).
#include "orbref.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

struct orbref_vocab {
    int k = 0, L = 0, scoring = 0, weighting = 0;
    std::vector<int32_t> parent, word_id;
    std::vector<double> weight;
    std::vector<uint8_t> desc;                     // [n][32]
    std::vector<std::vector<int32_t>> children;    // Node::children, TemplatedVocabulary.h:304
    int n_words = 0;
};

static void link_children(orbref_vocab *v) {
    const int n = (int)v->parent.size();
    v->children.assign(n, {});
    for (int i = 1; i < n; i++) v->children[v->parent[i]].push_back(i);   // ascending id = push_back order (:1392)
}

extern "C" {

int orbref_vocab_create(const orb_vocab_desc *d, orbref_vocab **out) {
    if (!d || !out || d->n_nodes < 1) return ORB_ERR_INVALID;
    orbref_vocab *v = new orbref_vocab;
    v->k = d->k; v->L = d->L; v->scoring = d->scoring; v->weighting = d->weighting;
    v->parent.assign(d->parent, d->parent + d->n_nodes);
    v->word_id.assign(d->word_id, d->word_id + d->n_nodes);
    v->weight.assign(d->weight, d->weight + d->n_nodes);
    v->desc.assign(d->descriptors, d->descriptors + (size_t)32 * d->n_nodes);
    for (int i = 1; i < d->n_nodes; i++)
        if (v->parent[i] < 0 || v->parent[i] >= d->n_nodes || v->parent[i] == i) { delete v; return ORB_ERR_INVALID; }
    link_children(v);
    for (int i = 1; i < d->n_nodes; i++) if (v->children[i].empty()) v->n_words++;
    *out = v;
    return ORB_OK;
}

// ref: TemplatedVocabulary::loadFromTextFile, TemplatedVocabulary.h:1338-1425.  Header "k L scoring weighting", then one
// line per node: parent, leaf flag, 32 descriptor bytes (FORB::fromString, FORB.cpp:123-138), weight.  The reference's
// loop runs `while(!f.eof())`, so after a final newline it runs once more on an empty read and uses an `int pid` that
// nothing has written (undefined behaviour; see include/orbgpu.h): that empty final read is not a node here.
int orbref_vocab_load_text(const char *path, orbref_vocab **out) {
    // C stdio on purpose: this library may be linked with a static libstdc++, whose iostreams must not be used next to
    // the process's shared one.  strtol / strtod read what the reference's stream extractions read.
    FILE *fp = std::fopen(path, "rb");
    if (!fp) return ORB_ERR_INVALID;
    std::string text;
    char buf[1 << 16];
    size_t got;
    while ((got = std::fread(buf, 1, sizeof buf, fp)) > 0) text.append(buf, got);
    std::fclose(fp);
    size_t pos = 0;
    bool eof = false;
    auto getline = [&](std::string &line) {   // std::getline: up to '\n'; eof is set when the end is hit while reading
        size_t nl = text.find('\n', pos);
        if (nl == std::string::npos) { line = text.substr(pos); pos = text.size(); eof = true; }
        else { line = text.substr(pos, nl - pos); pos = nl + 1; }
    };
    std::string line;
    getline(line);
    long hdr[4] = {0, 0, 0, 0};
    {
        const char *c = line.c_str(); char *q;
        for (int i = 0; i < 4; i++) { hdr[i] = std::strtol(c, &q, 10); if (q == c) { hdr[i] = 0; break; } c = q; }
    }
    const long k = hdr[0], L = hdr[1], n1 = hdr[2], n2 = hdr[3];
    if (k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) return ORB_ERR_INVALID;   // :1360-1364
    orbref_vocab *v = new orbref_vocab;
    v->k = (int)k; v->L = (int)L; v->scoring = (int)n1; v->weighting = (int)n2;
    v->parent.push_back(0); v->word_id.push_back(0); v->weight.push_back(0.0); v->desc.resize(32, 0);
    while (!eof) {
        getline(line);
        const char *c = line.c_str(); char *q;
        const long pid = std::strtol(c, &q, 10);
        if (q == c) {
            if (eof && line.find_first_not_of(" \t\r") == std::string::npos) break;   // the empty read after the last newline
            delete v;
            return ORB_ERR_INVALID;
        }
        c = q;
        bool ok = true;   // a failed extraction leaves the reference's stream failed: later reads yield nothing
        auto rd = [&]() -> long { if (!ok) return 0; long x = std::strtol(c, &q, 10); if (q == c) { ok = false; return 0; } c = q; return x; };
        const long leaf = rd();
        uint8_t d[32] = {0};
        for (int i = 0; i < 32; i++) { const long b = rd(); if (ok) d[i] = (uint8_t)b; }   // FORB::fromString, FORB.cpp:123-138
        double w = 0.0;
        if (ok) { w = std::strtod(c, &q); if (q == c) w = 0.0; }
        if (pid < 0 || pid > (long)v->parent.size()) { delete v; return ORB_ERR_INVALID; }   // the reference would index out of range
        v->parent.push_back((int32_t)pid);
        v->desc.insert(v->desc.end(), d, d + 32);
        v->weight.push_back(w);
        v->word_id.push_back(leaf > 0 ? v->n_words++ : 0);   // :1409-1416
    }
    link_children(v);
    *out = v;
    return ORB_OK;
}

void orbref_vocab_destroy(orbref_vocab *v) { delete v; }

int orbref_vocab_info(const orbref_vocab *v, int32_t *info) {
    info[0] = v->k; info[1] = v->L; info[2] = v->scoring; info[3] = v->weighting;
    info[4] = (int32_t)v->parent.size(); info[5] = v->n_words;
    return ORB_OK;
}

int orbref_vocab_nodes(const orbref_vocab *v, int32_t *parent, uint8_t *desc, double *weight, int32_t *word_id) {
    const size_t n = v->parent.size();
    if (parent) std::memcpy(parent, v->parent.data(), n * 4);
    if (desc) std::memcpy(desc, v->desc.data(), n * 32);
    if (weight) std::memcpy(weight, v->weight.data(), n * 8);
    if (word_id) std::memcpy(word_id, v->word_id.data(), n * 4);
    return ORB_OK;
}

}  // extern "C"

static int forb_distance(const uint8_t *a, const uint8_t *b) {   // FORB::distance, FORB.cpp:81-101 (same bit trick as ORBmatcher.cc:1650)
    int32_t pa[8], pb[8];
    std::memcpy(pa, a, 32); std::memcpy(pb, b, 32);
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        unsigned int x = pa[i] ^ pb[i];
        x = x - ((x >> 1) & 0x55555555);
        x = (x & 0x33333333) + ((x >> 2) & 0x33333333);
        dist += (((x + (x >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

// ref: transform(feature, word_id, weight, nid, levelsup), TemplatedVocabulary.h:1217-1255
static void descend(const orbref_vocab *v, const uint8_t *feat, int levelsup, int32_t *word, double *weight, int32_t *nid) {
    const int nid_level = v->L - levelsup;
    bool nid_set = false;
    if (nid_level <= 0) { *nid = 0; nid_set = true; }
    int final_id = 0, level = 0;
    do {
        ++level;
        const std::vector<int32_t> &nodes = v->children[final_id];
        final_id = nodes[0];
        double best = forb_distance(feat, &v->desc[(size_t)32 * final_id]);
        for (size_t c = 1; c < nodes.size(); c++) {
            double d = forb_distance(feat, &v->desc[(size_t)32 * nodes[c]]);
            if (d < best) { best = d; final_id = nodes[c]; }
        }
        if (level == nid_level) { *nid = final_id; nid_set = true; }
    } while (!v->children[final_id].empty());
    if (!nid_set) *nid = final_id;   // the reference leaves *nid unset here; include/orbgpu.h fixes it to the leaf
    *word = v->word_id[final_id];
    *weight = v->weight[final_id];
}

extern "C" {

int orbref_vocab_transform_features(const orbref_vocab *v, const uint8_t *desc, int n, int levelsup, int32_t *word,
                                    int32_t *node, double *weight) {
    if (v->n_words == 0 || v->children[0].empty()) return ORB_ERR_INVALID;
    for (int i = 0; i < n; i++) descend(v, desc + (size_t)32 * i, levelsup, &word[i], &weight[i], &node[i]);
    return ORB_OK;
}

// ref: transform(features, v, fv, levelsup), TemplatedVocabulary.h:1126-1193
int orbref_vocab_transform(const orbref_vocab *v, const uint8_t *desc, int n, int levelsup, int capacity, int32_t *bow_word,
                           double *bow_value, int32_t *n_bow, int32_t *fv_node, int32_t *fv_off, int32_t *fv_feat, int32_t *n_fv) {
    *n_bow = 0; *n_fv = 0; fv_off[0] = 0;
    if (v->n_words == 0) return ORB_OK;   // empty(), :1133-1136
    std::map<uint32_t, double> bow;
    std::map<uint32_t, std::vector<int32_t>> fv;
    const bool must = v->scoring != 5;                 // ScoringObject.h:74-89: only DotProductScoring skips normalisation
    const bool l2 = v->scoring == 1;
    const bool tf = v->weighting == 0 || v->weighting == 1;
    for (int i = 0; i < n; i++) {
        int32_t word, nid; double w;
        descend(v, desc + (size_t)32 * i, levelsup, &word, &w, &nid);
        if (w > 0) {
            auto it = bow.find((uint32_t)word);
            if (it == bow.end()) bow[(uint32_t)word] = w;          // addWeight inserts v / addIfNotExist inserts v
            else if (tf) it->second += w;                          // addWeight (BowVector.cpp:33-45); IDF/BINARY keep the first
            fv[(uint32_t)nid].push_back(i);
        }
    }
    if (tf && !bow.empty() && !must) {                 // :1163-1169
        const double nd = (double)bow.size();
        for (auto &e : bow) e.second /= nd;
    }
    if (must) {                                        // BowVector::normalize, BowVector.cpp:62-88
        double norm = 0.0;
        if (!l2) { for (auto &e : bow) norm += std::fabs(e.second); }
        else { for (auto &e : bow) norm += e.second * e.second; norm = std::sqrt(norm); }
        if (norm > 0.0) for (auto &e : bow) e.second /= norm;
    }
    if ((int)bow.size() > capacity || (int)fv.size() > capacity) return ORB_ERR_CAPACITY;
    int i = 0;
    for (auto &e : bow) { bow_word[i] = (int32_t)e.first; bow_value[i] = e.second; i++; }
    *n_bow = i;
    int g = 0, o = 0;
    for (auto &e : fv) {
        fv_node[g] = (int32_t)e.first; fv_off[g] = o;
        for (int32_t j : e.second) fv_feat[o++] = j;
        g++;
    }
    fv_off[g] = o;
    *n_fv = g;
    return ORB_OK;
}

}  // extern "C"
