// llama_logits_check.cpp -- TEST INFRASTRUCTURE (built into oracle/_ref next to the reference's libllama, never linked into the product).
//
// End-to-end parity of a WHOLE llama graph: the same GGUF is loaded twice by the reference's own loader -- once on its CPU backend
// (n_gpu_layers = 0), once with every layer offloaded to the backend found through GGML_BACKEND_PATH (this repository's plugin) -- a fixed token
// sequence is decoded as one prompt batch (prefill path: tensor-core GEMMs, tiled attention) and then token by token (decode path: GEMV groups,
// fused norm / attention kernels, whole-graph CUDA graph replay from the second token on), and the logits of every step are compared
// (NMSE, tolerance 2e-3 as the reference's own whole-model test, tests/test-backend-ops.cpp:6690-6692).
//   GGML_BACKEND_PATH=.../libggml-b200.so llama_logits_check model.gguf [n_prompt] [n_gen]
#include "llama.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "ggml.h"
#include "ggml-backend.h"
#include <string>
#include <cstring>

// LLAMA_LOGITS_TRACE=1: per-node statistics through the scheduler's eval callback (every node is computed on its own), to find the first node that differs
struct NodeStat { std::string name, op; double sum, asum; size_t bad, n; };
static bool trace_cb(struct ggml_tensor * t, bool ask, void * ud) {
    if (ask) return true;
    std::vector<NodeStat> * out = (std::vector<NodeStat> *) ud;
    if (t->type != GGML_TYPE_F32 || !ggml_is_contiguous(t)) return true;
    std::vector<float> v(ggml_nelements(t));
    ggml_backend_tensor_get(t, v.data(), 0, v.size() * sizeof(float));
    NodeStat st = { t->name, ggml_op_name(t->op), 0, 0, 0, v.size() };
    for (float x : v) { if (!std::isfinite(x)) ++st.bad; else { st.sum += x; st.asum += std::fabs(x); } }
    out->push_back(st);
    return true;
}

struct Run { llama_model * model; llama_context * ctx; int n_vocab; std::vector<NodeStat> * trace = nullptr; };

static Run open_model(const char * path, int ngl, int n_ctx) {
    llama_model_params mp = llama_model_default_params();
    mp.n_gpu_layers = ngl;
    mp.use_mmap = false;
    // LLAMA_LOGITS_SPLIT=row: weights through the backend's "ggml_backend_split_buffer_type" proc (llama-bench -sm row); =none: one device
    if (const char * sm = getenv("LLAMA_LOGITS_SPLIT")) {
        if (!strcmp(sm, "row")) mp.split_mode = LLAMA_SPLIT_MODE_ROW;
        else if (!strcmp(sm, "none")) mp.split_mode = LLAMA_SPLIT_MODE_NONE;
        else mp.split_mode = LLAMA_SPLIT_MODE_LAYER;
    }
    Run r = {};
    r.model = llama_model_load_from_file(path, mp);
    if (!r.model) { fprintf(stderr, "cannot load %s (ngl %d)\n", path, ngl); exit(2); }
    llama_context_params cp = llama_context_default_params();
    cp.n_ctx = n_ctx; cp.n_batch = n_ctx; cp.n_ubatch = 512; cp.n_threads = 8; cp.n_threads_batch = 8;
    // LLAMA_LOGITS_FA=0 / 1: flash attention off (what llama-bench defaults to: the MUL_MAT + SOFT_MAX attention path) / on; unset: the library's default (auto)
    if (const char * fa = getenv("LLAMA_LOGITS_FA")) cp.flash_attn_type = fa[0] == '1' ? LLAMA_FLASH_ATTN_TYPE_ENABLED : LLAMA_FLASH_ATTN_TYPE_DISABLED;
    if (getenv("LLAMA_LOGITS_TRACE")) { r.trace = new std::vector<NodeStat>(); cp.cb_eval = trace_cb; cp.cb_eval_user_data = r.trace; }
    r.ctx = llama_init_from_model(r.model, cp);
    if (!r.ctx) { fprintf(stderr, "cannot create a context (ngl %d)\n", ngl); exit(2); }
    r.n_vocab = llama_vocab_n_tokens(llama_model_get_vocab(r.model));
    return r;
}

static std::vector<float> step(Run & r, std::vector<llama_token> toks) {
    if (llama_decode(r.ctx, llama_batch_get_one(toks.data(), (int32_t) toks.size())) != 0) { fprintf(stderr, "llama_decode failed\n"); exit(3); }
    const float * l = llama_get_logits_ith(r.ctx, -1);
    return std::vector<float>(l, l + r.n_vocab);
}

int main(int argc, char ** argv) {
    if (argc < 2) { fprintf(stderr, "usage: llama_logits_check model.gguf [n_prompt] [n_gen]\n"); return 1; }
    const int n_prompt = argc > 2 ? atoi(argv[2]) : 40, n_gen = argc > 3 ? atoi(argv[3]) : 6;
    llama_backend_init();
    ggml_backend_load_all();
    Run cpu = open_model(argv[1], 0, 1024), gpu = open_model(argv[1], 99, 1024);
    std::vector<llama_token> prompt;
    uint32_t s = 12345;
    for (int i = 0; i < n_prompt; ++i) { s = s * 1664525u + 1013904223u; prompt.push_back((llama_token) (s % (uint32_t) cpu.n_vocab)); }
    int bad = 0;
    auto compare = [&](const char * what, const std::vector<float> & a, const std::vector<float> & b) {
        double num = 0, den = 0; bool finite = true;
        for (size_t i = 0; i < a.size(); ++i) { num += (double) (a[i] - b[i]) * (a[i] - b[i]); den += (double) b[i] * b[i]; finite = finite && std::isfinite(a[i]); }
        size_t bad_a = 0, bad_b = 0;
        for (size_t i = 0; i < a.size(); ++i) { bad_a += !std::isfinite(a[i]); bad_b += !std::isfinite(b[i]); }
        if (bad_a || bad_b) printf("  %-22s non-finite logits: backend %zu, cpu %zu of %zu; first values backend %g %g %g cpu %g %g %g\n", what, bad_a, bad_b, a.size(), a[0], a[1], a[2], b[0], b[1], b[2]);
        const double nmse = num / (den > 0 ? den : 1e-300);
        const bool ok = finite && den > 0 && nmse <= 2e-3;
        printf("  %-22s nmse %.3e %s\n", what, nmse, ok ? "ok" : "FAIL");
        bad += ok ? 0 : 1;
    };
    compare("prompt (prefill)", step(gpu, prompt), step(cpu, prompt));
    if (cpu.trace && gpu.trace) {
        int shown = 0;
        size_t jg = 0;
        for (size_t i = 0; i < cpu.trace->size() && shown < 10; ++i) {
            const NodeStat & b = (*cpu.trace)[i];
            size_t j = jg;
            while (j < gpu.trace->size() && (*gpu.trace)[j].name != b.name) ++j;      // the two schedules may hold different copy / view nodes
            if (j == gpu.trace->size()) continue;
            jg = j + 1;
            const NodeStat & a = (*gpu.trace)[j];
            const bool differ = a.bad != b.bad || std::fabs(a.asum - b.asum) > 1e-2 * (std::fabs(b.asum) + 1e-6);
            if (differ || i < 2) { printf("  node %3zu %-10s %-24s | backend sum %.5e |sum| %.5e bad %zu | cpu sum %.5e |sum| %.5e bad %zu %s\n", i, b.op.c_str(), b.name.c_str(), a.sum, a.asum, a.bad, b.sum, b.asum, b.bad, differ ? "<-- differs" : ""); shown += differ; }
        }
        cpu.trace->clear(); gpu.trace->clear();
    }
    for (int i = 0; i < n_gen; ++i) {
        s = s * 1664525u + 1013904223u;
        const std::vector<llama_token> t = { (llama_token) (s % (uint32_t) cpu.n_vocab) };
        char name[32]; snprintf(name, sizeof name, "token %d (decode)", i);
        compare(name, step(gpu, t), step(cpu, t));
    }
    printf("LLAMA_LOGITS_CHECK %s\n", bad ? "FAILED" : "OK");
    llama_free(cpu.ctx); llama_free(gpu.ctx); llama_model_free(cpu.model); llama_model_free(gpu.model);
    return bad ? 1 : 0;
}
