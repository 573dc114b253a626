// planes.cuh -- description of the B200 plane layout (see repack.cu header and DESIGN.md).
#pragma once
#include "common.cuh"

struct PlaneSpec {
    int n;               // number of planes
    int block_bytes;     // canonical block size
    int off[4];          // field offset inside the canonical block
    int len[4];          // field bytes per block
    int unit[4];         // copy unit used by the repack kernels (min(len, 16))
    int base_blocks[4];  // plane p starts at base_blocks[p] * (number of blocks) bytes
};

static inline bool plane_spec(int type, PlaneSpec * ps) {
    *ps = PlaneSpec{};
    auto set = [&](int bb, int n, const int * off, const int * len) {
        ps->n = n; ps->block_bytes = bb;
        int acc = 0;
        for (int i = 0; i < n; ++i) {
            ps->off[i] = off[i]; ps->len[i] = len[i]; ps->unit[i] = len[i] < 16 ? len[i] : 16;
            ps->base_blocks[i] = acc; acc += len[i];
        }
    };
    switch (type) {
        case B200_TYPE_Q4_0:   { const int o[] = {2, 0},             l[] = {16, 2};           set(18, 2, o, l);  return true; }
        case B200_TYPE_Q8_0:   { const int o[] = {2, 0},             l[] = {32, 2};           set(34, 2, o, l);  return true; }
        case B200_TYPE_Q4_K:   { const int o[] = {0},                l[] = {144};             set(144, 1, o, l); return true; }
        case B200_TYPE_Q5_K:   { const int o[] = {0},                l[] = {176};             set(176, 1, o, l); return true; }
        case B200_TYPE_Q6_K:   { const int o[] = {0, 128, 192, 208}, l[] = {128, 64, 16, 2};  set(210, 4, o, l); return true; }
        case B200_TYPE_IQ4_XS: { const int o[] = {8, 0},             l[] = {128, 8};          set(136, 2, o, l); return true; }
    }
    return false;
}
