// plugin_bench.cpp -- MEASUREMENT INFRASTRUCTURE (built into oracle/_ref next to the reference's libraries, never linked into the product).
//
// bench.py's `e2e_plugin` leg: the headline workload -- every quantised MUL_MAT of one Llama-3-8B Q4_K_M forward pass, 225 nodes -- built
// as ONE ggml_cgraph with the reference's public API and driven the way llama.cpp drives a backend:
//     ggml_backend_load_all (GGML_BACKEND_PATH -> libggml-b200.so) -> ggml_backend_alloc_ctx_tensors -> ggml_backend_tensor_set (weights: the
//     load-time repack hook; per step: the token's activations from host memory) -> ggml_backend_graph_compute -> ggml_backend_tensor_get
// (/root/reference/ggml/include/ggml-backend.h). Everything between set and get is the plugin's graph_compute (csrc/ggml_backend/ggml-b200.cpp).
// Weights are random-but-valid blocks of the model's per-tensor types (src/llama-quant.cpp:185-187, 302-303, 358-364), one random matrix per
// distinct (type, rows, k) uploaded into every tensor of that shape (distinct device tensors: 4.6 GB streamed per token).
//   GGML_BACKEND_PATH=.../libggml-b200.so plugin_bench [--n 1] [--steps 64] [--warmup 8]      -> one JSON line on stdout
#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <tuple>
#include <vector>

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static inline uint64_t rnd() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return rng_state; }
static inline float rnd01() { return (float) ((rnd() >> 40) * (1.0 / 16777216.0)); }

// uniform random payload, finite fp16 scale fields of the magnitude a N(0, 0.02^2) matrix would produce (as zyk-llama.cpp_b200/workloads.py)
static std::vector<uint8_t> random_blocks(ggml_type t, int64_t m, int64_t k) {
    const size_t row = ggml_row_size(t, k), n = row * (size_t) m;
    std::vector<uint8_t> w(n + 8);
    for (size_t i = 0; i < n; i += 8) { const uint64_t v = rnd(); memcpy(&w[i], &v, 8); }
    w.resize(n);
    const size_t bs = ggml_type_size(t), nb = n / bs;
    auto put = [&](size_t off, float v) { const ggml_fp16_t h = ggml_fp32_to_fp16(v); memcpy(&w[off], &h, 2); };
    for (size_t b = 0; b < nb; ++b) {
        const float mag = 0.25f + rnd01(), sign = (rnd() & 1) ? 1.0f : -1.0f;
        if (t == GGML_TYPE_Q4_K)      { put(b * bs, mag * 0.02f / 600.0f); put(b * bs + 2, (0.25f + rnd01()) * 0.02f / 40.0f); }
        else if (t == GGML_TYPE_Q6_K) { put(b * bs + 208, sign * mag * 0.02f / 2048.0f); }
        else { fprintf(stderr, "plugin_bench: type %s not handled\n", ggml_type_name(t)); exit(2); }
    }
    return w;
}

static bool use_more_bits(int i, int n) { return i < n / 8 || i >= 7 * n / 8 || (i - n / 8) % 3 == 2; }

int main(int argc, char ** argv) {
    int64_t n = 1; int steps = 64, warmup = 8;
    for (int i = 1; i + 1 < argc; i += 2) {
        if (!strcmp(argv[i], "--n")) n = atoll(argv[i + 1]);
        else if (!strcmp(argv[i], "--steps")) steps = atoi(argv[i + 1]);
        else if (!strcmp(argv[i], "--warmup")) warmup = atoi(argv[i + 1]);
    }
    ggml_backend_load_all();
    ggml_backend_dev_t gpu = nullptr;
    for (size_t i = 0; i < ggml_backend_dev_count(); ++i) {
        ggml_backend_dev_t d = ggml_backend_dev_get(i);
        if (ggml_backend_dev_type(d) == GGML_BACKEND_DEVICE_TYPE_GPU && !gpu) gpu = d;
    }
    if (!gpu) { fprintf(stderr, "plugin_bench: no GPU device (GGML_BACKEND_PATH set?)\n"); return 3; }
    ggml_backend_t be = ggml_backend_dev_init(gpu, nullptr);
    if (!be) { fprintf(stderr, "plugin_bench: backend init failed\n"); return 3; }

    const int n_layer = 32; const int64_t d = 4096, kv = 1024, ff = 14336, vocab = 128256;
    ggml_init_params ip = { ggml_tensor_overhead() * (size_t) (n_layer * 16 + 16) + ggml_graph_overhead_custom(1024, false), nullptr, true };
    ggml_context * ctx = ggml_init(ip);
    ggml_tensor * x0 = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, d, n), * attn = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, d, n), * act = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, ff, n);
    std::vector<ggml_tensor *> weights, outs;
    auto W = [&](ggml_type t, int64_t k, int64_t m) { ggml_tensor * w = ggml_new_tensor_2d(ctx, t, k, m); weights.push_back(w); return w; };
    ggml_tensor * x = x0;
    int n_nodes = 0;
    for (int l = 0; l < n_layer; ++l) {
        const ggml_type tv = use_more_bits(l, n_layer) ? GGML_TYPE_Q6_K : GGML_TYPE_Q4_K, td = tv;
        ggml_tensor * q = ggml_mul_mat(ctx, W(GGML_TYPE_Q4_K, d, d), x), * k = ggml_mul_mat(ctx, W(GGML_TYPE_Q4_K, d, kv), x), * v = ggml_mul_mat(ctx, W(tv, d, kv), x);
        ggml_tensor * x2 = ggml_mul_mat(ctx, W(GGML_TYPE_Q4_K, d, d), attn);
        ggml_tensor * g = ggml_mul_mat(ctx, W(GGML_TYPE_Q4_K, d, ff), x2), * u = ggml_mul_mat(ctx, W(GGML_TYPE_Q4_K, d, ff), x2);
        x = ggml_mul_mat(ctx, W(td, ff, d), act);
        for (ggml_tensor * t : { q, k, v, x2, g, u, x }) outs.push_back(t);
        n_nodes += 7;
    }
    ggml_tensor * logits = ggml_mul_mat(ctx, W(GGML_TYPE_Q6_K, d, vocab), x);
    outs.push_back(logits); ++n_nodes;
    ggml_cgraph * graph = ggml_new_graph_custom(ctx, 1024, false);
    for (ggml_tensor * t : outs) ggml_build_forward_expand(graph, t);
    ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(ctx, be);
    if (!buf) { fprintf(stderr, "plugin_bench: allocation failed\n"); return 2; }

    // weights through set_tensor (the plugin's load-time repack hook)
    std::map<std::tuple<int, int64_t, int64_t>, std::vector<uint8_t>> pool;
    size_t weight_bytes = 0;
    const auto t_load0 = std::chrono::steady_clock::now();
    for (ggml_tensor * w : weights) {
        auto key = std::make_tuple((int) w->type, w->ne[1], w->ne[0]);
        if (!pool.count(key)) pool[key] = random_blocks(w->type, w->ne[1], w->ne[0]);
        ggml_backend_tensor_set(w, pool[key].data(), 0, pool[key].size());
        weight_bytes += pool[key].size();
    }
    const double load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_load0).count();
    std::vector<float> hx((size_t) d * n), hattn((size_t) d * n), hact((size_t) ff * n), hlog((size_t) vocab);
    for (float & v : hx) v = rnd01() * 2 - 1;
    for (float & v : hattn) v = rnd01() * 2 - 1;
    for (float & v : hact) v = rnd01() * 2 - 1;
    ggml_backend_tensor_set(attn, hattn.data(), 0, hattn.size() * sizeof(float));
    ggml_backend_tensor_set(act, hact.data(), 0, hact.size() * sizeof(float));

    auto step = [&]() {
        ggml_backend_tensor_set(x0, hx.data(), 0, hx.size() * sizeof(float));                     // this step's input from host memory
        if (ggml_backend_graph_compute(be, graph) != GGML_STATUS_SUCCESS) { fprintf(stderr, "plugin_bench: graph_compute failed\n"); exit(4); }
        ggml_backend_tensor_get(logits, hlog.data(), (size_t) (n - 1) * vocab * sizeof(float), (size_t) vocab * sizeof(float));   // last token's logits
    };
    for (int i = 0; i < warmup; ++i) step();
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < steps; ++i) step();
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    double cs = 0; bool finite = true;
    for (float v : hlog) { cs += v; finite = finite && std::isfinite(v); }
    printf("{\"through\": \"ggml plugin: ggml_backend_tensor_set -> ggml_backend_graph_compute -> ggml_backend_tensor_get (reference API, GGML_BACKEND_PATH)\", "
           "\"device\": \"%s\", \"n\": %lld, \"nodes\": %d, \"steps\": %d, \"warmup\": %d, \"value\": %.3f, \"unit\": \"tokens/s\", \"ms_per_step\": %.4f, "
           "\"h2d_bytes_per_step\": %zu, \"d2h_bytes_per_step\": %zu, \"weight_bytes\": %zu, \"load_s\": %.2f, \"logits_finite\": %s, \"logits_checksum\": %.6e}\n",
           ggml_backend_dev_name(gpu), (long long) n, n_nodes, steps, warmup, (double) n * steps / secs, secs / steps * 1e3,
           hx.size() * sizeof(float), (size_t) vocab * sizeof(float), weight_bytes, load_s, finite ? "true" : "false", cs);
    ggml_backend_buffer_free(buf);
    ggml_free(ctx);
    ggml_backend_free(be);
    return finite ? 0 : 5;
}
