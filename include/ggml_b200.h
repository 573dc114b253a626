/* ggml_b200.h -- entry points of the ggml backend plugin libggml-b200.so.
 *
 * The plugin is the drop-in for the reference's backend plugin boundary
 * (/root/reference/ggml/src/ggml-backend-impl.h:11-251): it implements the five vtables
 * (reg / device / buffer_type / buffer / backend) on top of the C ABI in b200_mm.h, and is loaded by
 * the UNMODIFIED reference through
 *     GGML_BACKEND_PATH=/path/to/libggml-b200.so        (ggml/src/ggml-backend-reg.cpp:627-631)
 * or ggml_backend_load(path)                             (ggml/src/ggml-backend-reg.cpp:435-437),
 * both of which dlsym the two symbols below (ggml-backend-reg.cpp:262-306; typedefs
 * ggml-backend-impl.h:215-218). Everything else is reached through the returned registry object.
 *
 * Names obtainable through reg->iface.get_proc_address (ggml/include/ggml-backend.h:196-212):
 *     "ggml_backend_get_features"        -> ggml_backend_get_features_t
 *     anything else                      -> NULL ("ggml_backend_set_n_threads": no host threads on this path;
 *                                           "ggml_backend_split_buffer_type": not built yet -- row split across GPUs is
 *                                           available one process per GPU through b200_mul_mat_gather, b200_mm.h)
 *
 * Environment switch read by the plugin:
 *     GGML_B200_REPACK=0     keep weights in canonical ggml layout (default: repack whole-tensor writes)
 */
#ifndef GGML_B200_H
#define GGML_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define GGML_B200_API __attribute__((visibility("default")))

struct ggml_backend_reg;

/* ggml_backend_init_t (ggml-backend-impl.h:216): the registry object, static lifetime */
GGML_B200_API struct ggml_backend_reg * ggml_backend_init(void);
/* ggml_backend_score_t (ggml-backend-impl.h:218): 0 = no sm_100 device visible, refuse the load */
GGML_B200_API int ggml_backend_score(void);

#ifdef __cplusplus
}
#endif
#endif /* GGML_B200_H */
