// oracle/llama_quantize_types.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Runs the reference's own quantiser (llama_model_quantize, include/llama.h; per-tensor type selection in
// src/llama-quant.cpp:178-440 llama_tensor_get_type) on an f32 GGUF and prints the type it gave every tensor, one
// "name type" line each. tests/test_quant_recipe.py uses it to pin the per-tensor recipe of zyk-llama.cpp_b200/workloads.py
// (which tools/make_gguf.py and bench.py build their random-init models from) against what llama-quantize really writes
// for the BASELINE.json models.
//
//     llama_quantize_types in-f32.gguf out.gguf <llama_ftype number> [threads]
#include "ggml.h"
#include "gguf.h"
#include "llama.h"

#include <cstdio>
#include <cstdlib>

int main(int argc, char ** argv) {
    if (argc < 4) {
        fprintf(stderr, "usage: %s in-f32.gguf out.gguf ftype [threads]\n", argv[0]);
        return 2;
    }
    llama_log_set([](ggml_log_level level, const char * text, void *) { if (level >= GGML_LOG_LEVEL_ERROR) fputs(text, stderr); }, nullptr);
    llama_model_quantize_params params = llama_model_quantize_default_params();
    params.ftype   = (llama_ftype) atoi(argv[3]);
    params.nthread = argc > 4 ? atoi(argv[4]) : 8;
    if (llama_model_quantize(argv[1], argv[2], &params) != 0) {
        fprintf(stderr, "llama_model_quantize failed\n");
        return 1;
    }
    gguf_init_params ip = { /*no_alloc*/ true, /*ctx*/ nullptr };
    gguf_context * g = gguf_init_from_file(argv[2], ip);
    if (!g) {
        fprintf(stderr, "cannot read %s back\n", argv[2]);
        return 1;
    }
    const int64_t n = gguf_get_n_tensors(g);
    for (int64_t i = 0; i < n; ++i) {
        printf("%s %s\n", gguf_get_tensor_name(g, i), ggml_type_name(gguf_get_tensor_type(g, i)));
    }
    gguf_free(g);
    return 0;
}
