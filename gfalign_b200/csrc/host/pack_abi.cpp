// gfalign_b200/csrc/host/pack_abi.cpp — C ABI over the text packers (include/gfalign_b200.h).
#include <cstdio>
#include <cstring>
#include <new>

#include "host.h"

struct gfe_text {
    gfh::Graph graph;
    gfh::Gaf gaf;
    bool have_gaf = false;
};

extern "C" {

int gfe_text_open(const char *gfa_path, const char *gaf_path, int threads, gfe_text **out,
                  char *err, size_t err_len) {
    auto fail = [&](const std::string &m, int rc) {
        if (err && err_len) snprintf(err, err_len, "%s", m.c_str());
        return rc;
    };
    if (!gfa_path || !out) return fail("gfe_text_open: NULL argument", GFE_ERR_ARG);
    *out = nullptr;
    gfe_text *t = new (std::nothrow) gfe_text();
    if (!t) return fail("out of memory", GFE_ERR_NOMEM);
    std::string msg;
    try {
        if (!t->graph.load(gfa_path, msg)) {
            delete t;
            return fail(msg, GFE_ERR_ARG);
        }
        if (gaf_path) {
            if (threads <= 0) threads = gfh::hardware_threads();
            if (!t->gaf.load(gaf_path, t->graph.names, 0, threads, false, msg)) {
                delete t;
                return fail(msg, GFE_ERR_ARG);
            }
            t->have_gaf = true;
        }
    } catch (const std::bad_alloc &) {
        delete t;
        return fail("out of memory", GFE_ERR_NOMEM);
    }
    *out = t;
    return GFE_OK;
}

void gfe_text_close(gfe_text *t) { delete t; }

int gfe_text_get(const gfe_text *t, gfe_text_view *v) {
    if (!t || !v) return GFE_ERR_ARG;
    v->n_nodes = t->graph.names.size();
    v->n_links = (uint32_t)t->graph.links.size();
    v->edge_from = t->graph.edge_from.data();
    v->edge_to = t->graph.edge_to.data();
    v->graph = t->graph.view();
    v->A = t->have_gaf ? t->gaf.A : 0;
    v->offsets = t->have_gaf ? t->gaf.offsets.data() : nullptr;
    v->steps = t->have_gaf ? t->gaf.steps.data() : nullptr;
    return GFE_OK;
}

const char *gfe_text_node_name(const gfe_text *t, uint32_t node) {
    if (!t || node >= t->graph.names.size()) return nullptr;
    return t->graph.names.name(node).c_str();
}

int64_t gfe_text_node_id(const gfe_text *t, const char *name) {
    if (!t || !name) return -1;
    return t->graph.names.find(name);
}

const char *gfe_text_qname(const gfe_text *t, uint64_t a, uint32_t *len) {
    if (!t || !t->have_gaf || a >= t->gaf.A) return nullptr;
    if (len) *len = t->gaf.qname_len[a];
    return t->gaf.bytes.data + t->gaf.line_off[a];
}

}  // extern "C"
