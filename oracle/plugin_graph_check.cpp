// plugin_graph_check.cpp -- TEST INFRASTRUCTURE (built into oracle/_ref, never linked into the product).
//
// The reference's test-backend-ops only builds one-node graphs. This program drives the plugin the way llama.cpp's decode
// graph does: SEVERAL dependent MUL_MAT nodes in one ggml_cgraph,
//     {wq, wk, wv}(x) -> wo(q) -> {ffn_gate, ffn_up}(o) -> ffn_down(gate)
// through the reference's own public API (ggml_backend_load_all -> GGML_BACKEND_PATH, ggml_backend_graph_compute,
// /root/reference/ggml/include/ggml-backend.h), once on the reference CPU backend and once on the B200 backend, and compares
// every output with the tolerance of test-backend-ops (NMSE <= 5e-4, tests/test-backend-ops.cpp:3716-3718).
// It exercises the plugin's grouping (b200_mul_mat_group) and decode-chain (b200_chain_*) paths of graph_compute.
//   GGML_BACKEND_PATH=.../libggml-b200.so plugin_graph_check [n]        (n = activation rows, default 1)
#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

struct Model {
    ggml_context * ctx = nullptr;
    ggml_backend_buffer_t buf = nullptr;
    ggml_tensor * x = nullptr;
    std::vector<ggml_tensor *> weights, outs;
    ggml_cgraph * graph = nullptr;
};

static ggml_tensor * weight(ggml_context * ctx, ggml_type t, int64_t k, int64_t m, const char * name, std::vector<ggml_tensor *> & all) {
    ggml_tensor * w = ggml_new_tensor_2d(ctx, t, k, m);
    ggml_set_name(w, name);
    all.push_back(w);
    return w;
}

static Model build(ggml_backend_t backend, int64_t n, const std::vector<std::vector<uint8_t>> & wdata, const std::vector<float> & xdata) {
    Model M;
    ggml_init_params ip = { ggml_tensor_overhead() * 64 + ggml_graph_overhead(), nullptr, true };
    M.ctx = ggml_init(ip);
    const int64_t d = 4096, kv = 1024, ff = 14336;
    M.x = ggml_new_tensor_2d(M.ctx, GGML_TYPE_F32, d, n);
    ggml_tensor * wq = weight(M.ctx, GGML_TYPE_Q4_K, d, d, "wq", M.weights);
    ggml_tensor * wk = weight(M.ctx, GGML_TYPE_Q4_K, d, kv, "wk", M.weights);
    ggml_tensor * wv = weight(M.ctx, GGML_TYPE_Q6_K, d, kv, "wv", M.weights);
    ggml_tensor * wo = weight(M.ctx, GGML_TYPE_Q4_K, d, d, "wo", M.weights);
    ggml_tensor * wg = weight(M.ctx, GGML_TYPE_Q4_K, d, ff, "ffn_gate", M.weights);
    ggml_tensor * wu = weight(M.ctx, GGML_TYPE_Q4_K, d, ff, "ffn_up", M.weights);
    ggml_tensor * wd = weight(M.ctx, GGML_TYPE_Q6_K, ff, d, "ffn_down", M.weights);
    ggml_tensor * q = ggml_mul_mat(M.ctx, wq, M.x);
    ggml_tensor * k = ggml_mul_mat(M.ctx, wk, M.x);
    ggml_tensor * v = ggml_mul_mat(M.ctx, wv, M.x);
    ggml_tensor * o = ggml_mul_mat(M.ctx, wo, q);
    ggml_tensor * g = ggml_mul_mat(M.ctx, wg, o);
    ggml_tensor * u = ggml_mul_mat(M.ctx, wu, o);
    ggml_tensor * y = ggml_mul_mat(M.ctx, wd, g);
    M.outs = { q, k, v, o, g, u, y };
    M.graph = ggml_new_graph(M.ctx);
    for (ggml_tensor * t : M.outs) ggml_build_forward_expand(M.graph, t);
    M.buf = ggml_backend_alloc_ctx_tensors(M.ctx, backend);
    if (!M.buf) { fprintf(stderr, "alloc failed on %s\n", ggml_backend_name(backend)); exit(2); }
    for (size_t i = 0; i < M.weights.size(); ++i) ggml_backend_tensor_set(M.weights[i], wdata[i].data(), 0, wdata[i].size());
    ggml_backend_tensor_set(M.x, xdata.data(), 0, xdata.size() * sizeof(float));
    return M;
}

static std::vector<float> fetch(ggml_tensor * t) {
    std::vector<float> v(ggml_nelements(t));
    ggml_backend_tensor_get(t, v.data(), 0, v.size() * sizeof(float));
    return v;
}

int main(int argc, char ** argv) {
    const int64_t n = argc > 1 ? atoll(argv[1]) : 1;
    ggml_backend_load_all();
    ggml_backend_dev_t gpu = nullptr, cpu = nullptr;
    for (size_t i = 0; i < ggml_backend_dev_count(); ++i) {
        ggml_backend_dev_t d = ggml_backend_dev_get(i);
        if (ggml_backend_dev_type(d) == GGML_BACKEND_DEVICE_TYPE_GPU && !gpu) gpu = d;
        if (ggml_backend_dev_type(d) == GGML_BACKEND_DEVICE_TYPE_CPU && !cpu) cpu = d;
    }
    if (!gpu || !cpu) { fprintf(stderr, "need a CPU and a GPU device (GGML_BACKEND_PATH set?)\n"); return 3; }
    ggml_backend_t bg = ggml_backend_dev_init(gpu, nullptr), bc = ggml_backend_dev_init(cpu, nullptr);
    printf("devices: %s | %s, n = %lld\n", ggml_backend_dev_name(cpu), ggml_backend_dev_name(gpu), (long long) n);

    // weights: N(0, 0.02^2) quantised by the reference's own quantisers
    std::mt19937 rng(1234);
    std::normal_distribution<float> nd(0.0f, 1.0f);
    const int64_t d = 4096, kv = 1024, ff = 14336;
    struct Spec { ggml_type t; int64_t k, m; } specs[] = { { GGML_TYPE_Q4_K, d, d }, { GGML_TYPE_Q4_K, d, kv }, { GGML_TYPE_Q6_K, d, kv }, { GGML_TYPE_Q4_K, d, d },
                                                             { GGML_TYPE_Q4_K, d, ff }, { GGML_TYPE_Q4_K, d, ff }, { GGML_TYPE_Q6_K, ff, d } };
    std::vector<std::vector<uint8_t>> wdata;
    for (const Spec & s : specs) {
        std::vector<float> f((size_t) s.k * s.m);
        for (float & v : f) v = 0.02f * nd(rng);
        std::vector<uint8_t> qd(ggml_row_size(s.t, s.k) * s.m);
        ggml_quantize_chunk(s.t, f.data(), qd.data(), 0, s.m, s.k, nullptr);
        wdata.push_back(std::move(qd));
    }
    std::vector<float> xdata((size_t) d * n);
    for (float & v : xdata) v = nd(rng);

    Model mc = build(bc, n, wdata, xdata), mg = build(bg, n, wdata, xdata);
    if (ggml_backend_graph_compute(bc, mc.graph) != GGML_STATUS_SUCCESS) { fprintf(stderr, "CPU graph failed\n"); return 4; }
    int bad = 0;
    for (int pass = 0; pass < 3; ++pass) {                 // pass 1, 2: the cached chain / a changed input
        if (pass == 2) {
            for (float & v : xdata) v = nd(rng);
            ggml_backend_tensor_set(mc.x, xdata.data(), 0, xdata.size() * sizeof(float));
            ggml_backend_tensor_set(mg.x, xdata.data(), 0, xdata.size() * sizeof(float));
            if (ggml_backend_graph_compute(bc, mc.graph) != GGML_STATUS_SUCCESS) return 4;
        }
        if (ggml_backend_graph_compute(bg, mg.graph) != GGML_STATUS_SUCCESS) { fprintf(stderr, "GPU graph failed\n"); return 5; }
        for (size_t i = 0; i < mc.outs.size(); ++i) {
            const std::vector<float> a = fetch(mg.outs[i]), b = fetch(mc.outs[i]);
            double num = 0, den = 0; bool finite = true;
            for (size_t j = 0; j < a.size(); ++j) { num += (double) (a[j] - b[j]) * (a[j] - b[j]); den += (double) b[j] * b[j]; finite = finite && std::isfinite(a[j]); }
            const double nmse = num / (den > 0 ? den : 1e-300);
            const bool ok = finite && den > 0 && nmse <= 5e-4;
            if (!ok || pass == 0) printf("  pass %d %-9s nmse %.3e %s\n", pass, mc.weights[i]->name, nmse, ok ? "ok" : "FAIL");
            bad += ok ? 0 : 1;
        }
    }
    printf("PLUGIN_GRAPH_CHECK %s\n", bad ? "FAILED" : "OK");
    return bad ? 1 : 0;
}
