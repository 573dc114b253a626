"""zyk-llama.cpp_b200 -- B200-native quantised MUL_MAT / MUL_MAT_ID path (host-side bindings).

The product is two shared libraries built in-tree by ``csrc/Makefile``:

* ``lib/libb200mm.so``    -- hand-written sm_100a kernels behind the C ABI of ``include/b200_mm.h``
* ``lib/libggml-b200.so`` -- the ggml backend plugin (``ggml_backend_init``) the reference loads
  through ``GGML_BACKEND_PATH`` (``ggml/src/ggml-backend-reg.cpp:627-631``)

This Python package is only the ctypes view of the C ABI used by tests/ and bench.py; torch is
used for device memory and ``torch.distributed`` and nothing else. There is no CPU fallback: if
the extension is missing the import fails loudly.

The directory name is not a Python identifier; load it with ``__graft_entry__.load_pkg()`` which
registers it as module ``zyk_llama_cpp_b200``.
"""
from .binding import (  # noqa: F401
    B200Error,
    LIB_PATH,
    PLUGIN_PATH,
    lib,
    Types,
    LAYOUT_CANONICAL,
    LAYOUT_PLANES,
    Q8_REF,
    Q8_X86,
    MM_X_PREQUANTISED,
    MM_PDL,
    row_size,
    act_type,
    repack_supported,
    quantize_rows,
    repack,
    unrepack,
    mul_mat,
    mul_mat_id,
    mul_mat_group,
    MulMatArgs,
    DecodeChain,
    mul_mat_workspace,
    Stream,
    Event,
    Graph,
    launch_count,
    row_split,
)
