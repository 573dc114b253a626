// gemv_chain_xreg.cuh -- the decode chain kernel with the ACTIVATIONS STATIONARY IN REGISTERS (included by gemv_chain.cu, inside
// its anonymous namespace; n = 1, Q4_K / swizzled Q6_K nodes; everything else keeps gemv_chain_kernel).
//
// Why: in gemv_chain_kernel every 256-block of weights (144 B) costs 256 B of activation loads from shared memory on top of the
// weights themselves; the consumers drain the ring at about the speed HBM fills it (shared memory ~90 % busy), so the HBM time lost at
// a node boundary is never won back. A lane of this kernel works on the SAME 256-element unit of X for every row of a node (one
// block per lane, rows side by side when a row has fewer than 32 units, units 32 s .. 32 s + 31 for warp class s when it has more), so
// the unit's quantised X -- 256 int8 + the sums of 16 and of 32 + d, 77 registers -- is loaded once per node and shared memory carries
// only the weight stream: 9 (Q4_K) / 14 (Q6_K) LDS.128 per block instead of 27 / 33. The consumers become about twice as fast as
// HBM, the ring is empty when a node ends, and the boundary (last rows arrive, X is re-quantised) overlaps the refill of the ring.
//   * 8 consumer warps + one producer thread per CTA, ONE ring of NS stages (default 6) per CTA;
//   * the quantised X is double buffered in shared memory (node parity), so a node costs ONE CTA barrier: quantise -> barrier -> every
//     lane loads its unit; nobody touches the buffer again, the next node writes the other one;
//   * the tasks of consecutive items are dealt to the warps round robin ACROSS items (`rot` in the item descriptor), a warp with no
//     task in an item releases the stage at once and moves on, so no warp idles because an item has seven row pairs and the CTA eight warps;
//   * rows longer than 32 units: warp class s = warp % segs owns units [32 s, 32 s + 32); the partial sums meet in a per-stage buffer,
//     every warp arrives on the stage's `part` mbarrier and only the last class waits for it and stores Y (no CTA-wide barrier).
// Integer arithmetic, fp32 expressions and summation order are those of gemv_chain_kernel: results are bit-identical to it.

constexpr int CX_MAX_NS = 8;
constexpr int CX_HDR = 1024;          // full[8] @0, empty[8] @64, part[8] @128, item descriptors 8 x 48 @256
constexpr int CX_PART_FLOATS = 64;    // partial sums per stage (rows x segs of an item whose rows are cut)

struct XRegs { uint4 q[16]; uint4 b16[2]; uint4 b32; float d; };

__device__ __forceinline__ void xregs_load(XRegs & x, uint32_t xb) {
#pragma unroll
    for (int ch = 0; ch < 16; ++ch) x.q[ch] = lds16(xb + (uint32_t) (xpos(ch) * 16));
    x.b16[0] = lds16(xb + XU_BS16); x.b16[1] = lds16(xb + XU_BS16 + 16);
    x.b32 = lds16(xb + XU_BS32);
    x.d = ldsf(xb + XU_D);
}

// unit_q4k with X in registers (same arithmetic, same order)
__device__ __forceinline__ float unit_q4k_r(uint32_t wb, const XRegs & x) {
    const uint4 hdr = lds16(wb);
    uint4 q[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) q[i] = lds16(wb + 16 + 16 * i);
    const uint32_t sc03 = hdr.y & 0x3f3f3f3fu, mn03 = hdr.z & 0x3f3f3f3fu;
    const uint32_t sc47 = (hdr.w & 0x0f0f0f0fu) | ((hdr.y >> 2) & 0x30303030u);
    const uint32_t mn47 = ((hdr.w >> 4) & 0x0f0f0f0fu) | ((hdr.z >> 2) & 0x30303030u);
    const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&hdr.x));
    int isum = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint4 la = and4(q[2 * j], 0x0f0f0f0fu), lb = and4(q[2 * j + 1], 0x0f0f0f0fu);
        const uint4 ha = and4(q[2 * j], 0xf0f0f0f0u), hb = and4(q[2 * j + 1], 0xf0f0f0f0u);
        const uint32_t scw = j < 2 ? sc03 : sc47;
        const int s0 = (int) ((scw >> (16 * (j & 1))) & 0xffu), s1 = (int) ((scw >> (16 * (j & 1) + 8)) & 0xffu);
        const int dlo = dot16(la, x.q[4 * j]) + dot16(lb, x.q[4 * j + 1]);
        const int dhi = dot16_us(ha, x.q[4 * j + 2]) + dot16_us(hb, x.q[4 * j + 3]);
        isum += s0 * dlo + s1 * (dhi >> 4);
    }
    int msum = __dp2a_lo((int) x.b32.x, (int) mn03, 0);
    msum = __dp2a_hi((int) x.b32.y, (int) mn03, msum);
    msum = __dp2a_lo((int) x.b32.z, (int) mn47, msum);
    msum = __dp2a_hi((int) x.b32.w, (int) mn47, msum);
    return x.d * (dm.x * (float) isum - dm.y * (float) msum);
}

// unit_q6k with X in registers (swizzled stage layout, see unit_q6k)
__device__ __forceinline__ float unit_q6k_r(uint32_t stage, int nunits, int u, const XRegs & x) {
    const uint32_t qh0 = stage + (uint32_t) q6sw_qh0(nunits);
    const uint32_t qlp = (stage + (uint32_t) (u * 128)) ^ (uint32_t) ((u & 7) << 4);
    const uint32_t qhp = (qh0 + (uint32_t) (u * 64)) ^ (uint32_t) (((u >> 1) & 3) << 4);
    const uint4 scw = lds16(qh0 + (uint32_t) (nunits * 64 + u * 16));
    const float dw = half_bits_to_float(lds2(qh0 + (uint32_t) (nunits * 80 + u * 2)));
    uint4 ql[8], qh[4];
#pragma unroll
    for (int i = 0; i < 8; ++i) ql[i] = lds16(qlp ^ (uint32_t) (i << 4));
#pragma unroll
    for (int i = 0; i < 4; ++i) qh[i] = lds16(qhp ^ (uint32_t) (i << 4));
    const uint32_t sc[4] = {scw.x, scw.y, scw.z, scw.w};
    int m = __dp2a_lo((int) x.b16[0].x, (int) sc[0], 0);
    m = __dp2a_hi((int) x.b16[0].y, (int) sc[0], m);
    m = __dp2a_lo((int) x.b16[0].z, (int) sc[1], m);
    m = __dp2a_hi((int) x.b16[0].w, (int) sc[1], m);
    m = __dp2a_lo((int) x.b16[1].x, (int) sc[2], m);
    m = __dp2a_hi((int) x.b16[1].y, (int) sc[2], m);
    m = __dp2a_lo((int) x.b16[1].z, (int) sc[3], m);
    m = __dp2a_hi((int) x.b16[1].w, (int) sc[3], m);
    int isum = -32 * m;
#pragma unroll
    for (int n = 0; n < 2; ++n) {
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const uint4 qlv = ql[4 * n + t];
            const uint4 qhv = qh[2 * n + (t & 1)];
            constexpr uint32_t M3 = 0x03030303u;
            const int sh = t & 2;
            const uint4 lo = and4(qlv, 0x0f0f0f0fu), hi = and4(qlv, 0xf0f0f0f0u);
            const uint4 ha = and4(qhv, M3 << sh), hb = and4(qhv, M3 << (sh + 4));
            const int is = 8 * n + t;
            const int sc0 = sext8(sc[is >> 2], is & 3), sc1 = sext8(sc[(is + 4) >> 2], is & 3);
            const int s0 = dot16(lo, x.q[is]) + (dot16(ha, x.q[is]) << (4 - sh));
            const int s1 = (dot16_us(hi, x.q[is + 4]) >> 4) + (dot16_us(hb, x.q[is + 4]) >> sh);
            isum += sc0 * s0 + sc1 * s1;
        }
    }
    return (x.d * dw) * (float) isum;
}

template <int NW>
__global__ void __launch_bounds__((NW + 1) * 32, 1) gemv_chainx_kernel(const ChainArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int NN = a.n_nodes, NS = a.ns;

    const uint32_t base = smem_u32(smem_raw);
    auto full_bar  = [&](int st) { return base + (uint32_t) (st * 8); };
    auto empty_bar = [&](int st) { return base + 64u + (uint32_t) (st * 8); };
    auto part_bar  = [&](int st) { return base + 128u + (uint32_t) (st * 8); };
    auto desc_at   = [&](int st) { return base + 256u + (uint32_t) (st * 48); };
    const uint32_t xs0 = base + CX_HDR;
    const uint32_t part0 = xs0 + 2u * (uint32_t) a.x_bytes;
    const uint32_t ring0 = (part0 + (uint32_t) (NS * CX_PART_FLOATS * 4) + 1023u) & ~1023u;

    if (tid == 0) {
        for (int i = 0; i < NS; ++i) { mbar_init(full_bar(i), 1); mbar_init(empty_bar(i), NW); mbar_init(part_bar(i), NW); }
        fence_barrier_init();
    }
    __syncthreads();

    // ---- producer: one thread keeps the ring full; the item list of the CTA is node-ordered, so the copies run ahead of the node
    //      the consumers are working on for as long as there are free stages
    if (warp == NW) {
        if (lane != 0) return;
        const uint64_t policy = policy_evict_first();
        const int * off = a.node_off + (long long) blockIdx.x * (NN + 1);
        int nxt = off[0];
        const int end = off[NN];
        int st = 0; uint32_t ph = 0;
        ChainItemD item = {};
        if (nxt < end) item = a.items[nxt];
        while (nxt < end) {
            mbar_wait(empty_bar(st), ph ^ 1u);
            const uint32_t dst = ring0 + (uint32_t) st * (uint32_t) a.stage_bytes, bar = full_bar(st), dsc = desc_at(st);
            const unsigned long long yp = (unsigned long long) item.y, ld = (unsigned long long) item.ldy;
            asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" :: "r"(dsc), "r"((uint32_t) yp), "r"((uint32_t) (yp >> 32)), "r"((uint32_t) ld), "r"((uint32_t) (ld >> 32)) : "memory");
            asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" :: "r"(dsc + 16), "r"((uint32_t) item.type), "r"((uint32_t) item.nrows), "r"((uint32_t) item.segs), "r"((uint32_t) item.S) : "memory");
            { const unsigned long long tg = (unsigned long long) item.tag; sts8(dsc + 32, (uint32_t) tg, (uint32_t) (tg >> 32)); }
            const uint32_t total = item.bytes[0] + item.bytes[1] + item.bytes[2] + item.bytes[3];
            mbar_arrive_expect_tx(bar, total);
            if (item.type == CH_T_Q6K_SW) {
                const int nr = item.nrows, u0 = (int) (item.pad & 0xffffffffll), R = (int) (item.pad >> 32);
                const int upr_i = (int) (item.bytes[0] >> 7) / nr;
                const uint32_t qh0 = dst + (uint32_t) q6sw_qh0(nr * upr_i);
                int r = 0;
                for (; r + R <= nr; r += R) {
                    tma_load_2d_hint(dst + (uint32_t) (r * upr_i * 128), item.src[0], 0, u0 + r * upr_i, bar, policy);
                    tma_load_2d_hint(qh0 + (uint32_t) (r * upr_i * 64), item.src[1], 0, u0 + r * upr_i, bar, policy);
                }
                for (; r < nr; ++r) {
                    tma_load_2d_hint(dst + (uint32_t) (r * upr_i * 128), item.src[0] + 2 * sizeof(CUtensorMap), 0, u0 + r * upr_i, bar, policy);
                    tma_load_2d_hint(qh0 + (uint32_t) (r * upr_i * 64), item.src[1] + 2 * sizeof(CUtensorMap), 0, u0 + r * upr_i, bar, policy);
                }
                bulk_g2s(qh0 + item.bytes[1], item.src[2], item.bytes[2], bar, policy);
                bulk_g2s(qh0 + item.bytes[1] + item.bytes[2], item.src[3], item.bytes[3], bar, policy);
            } else {
                uint32_t o = 0;
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    if (item.bytes[p]) bulk_g2s(dst + o, item.src[p], item.bytes[p], bar, policy);
                    o += item.bytes[p];
                }
            }
            ++nxt;
            if (nxt < end) item = a.items[nxt];
            if (++st == NS) { st = 0; ph ^= 1u; }
        }
        return;
    }

    // ---- consumers ---------------------------------------------------------------------------------------------------------
    auto cons_bar_x = [] { asm volatile("bar.sync 3, %0;" :: "n"(NW * 32) : "memory"); };
    const int * off = a.node_off + (long long) blockIdx.x * (NN + 1);
    long long * tl = a.timeline ? a.timeline + ((long long) blockIdx.x * NN) * 8 : nullptr;

    const unsigned long long launch = *a.launches;
    const unsigned epoch0 = (unsigned) (launch * (unsigned long long) (NN + 1));
    const unsigned long long arrivals0 = launch * a.per_launch;
    int st = 0; uint32_t ph = 0, pmask = 0;
    ChainNodeD nd = a.nodes[0];
    int cnt = off[1] - off[0];
    for (int n = 0; n < NN; ++n) {
        ChainNodeD nd_next = nd; int cnt_next = 0;
        if (n + 1 < NN) { nd_next = a.nodes[n + 1]; cnt_next = off[n + 2] - off[n + 1]; }
        if (tl && tid == 0) tl[n * 8 + 0] = clock64();
        if (nd.mode != NB_NONE && nd.mode != NB_TAGGED && !(a.debug & 1)) {
            if (tid == 0) {
                const bool local = nd.mode == NB_RELAXED_LOCAL;
                const unsigned long long want = local ? launch * a.per_launch_local + (unsigned long long) nd.want : arrivals0 + (unsigned long long) nd.want * (unsigned) a.world;
                const unsigned long long * cp = local ? a.counter_local : a.counter;
                unsigned long long v, t0 = 0;
                for (unsigned spin = 0;; ++spin) {
                    if (a.world > 1 && !local) asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(cp) : "memory");
                    else                       asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(cp) : "memory");
                    if (v >= want) {
                        if (nd.mode == NB_FENCED) { if (a.world > 1) asm volatile("fence.acq_rel.sys;" ::: "memory"); else asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
                        break;
                    }
                    if ((spin & 4095u) == 4095u) {
                        unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                        if (t0 == 0) t0 = t; else if (t - t0 > 4000000000ull) __trap();
                    }
                }
            }
            cons_bar_x();
        }
        if (tl && tid == 0) tl[n * 8 + 1] = clock64();
        const uint32_t xs = xs0 + (uint32_t) (n & 1) * (uint32_t) a.x_bytes;
        if (!(a.debug & 4)) chain_quantise_x<1, NW>(nd, xs, warp, lane, a.tag_stride, epoch0 + (unsigned) n, a.world > 1);
        cons_bar_x();

        // the lane's unit of X for this node
        const int upr = nd.k >> 8;
        int rp = 1, sub = 0, b, segs = 1, seg = 0, idx = warp, per = NW;
        if (upr <= 32) { rp = 32 / upr; sub = lane / upr; b = lane - sub * upr; }
        else { segs = (upr + 31) >> 5; seg = warp % segs; idx = warp / segs; per = NW / segs; b = seg * 32 + lane; }
        const bool active = upr <= 32 ? sub < rp : b < upr;
        XRegs x;
        xregs_load(x, xs + (uint32_t) ((active ? b : 0) * XU_BYTES));
        if (tl && tid == 0) tl[n * 8 + 2] = clock64();

        long long c_wait = 0, c_loop = tl ? clock64() : 0;
        for (int i = 0; i < cnt; ++i) {
            const uint32_t stage = ring0 + (uint32_t) st * (uint32_t) a.stage_bytes;
            const long long w0 = tl ? clock64() : 0;
            mbar_wait(full_bar(st), ph);
            if (tl) c_wait += clock64() - w0;
            const uint4 d0 = lds16(desc_at(st));
            const uint4 d1 = lds16(desc_at(st) + 16);
            const uint2 d2 = lds8(desc_at(st) + 32);
            YOut out;
            out.y = (float *) (((unsigned long long) d0.y << 32) | d0.x);
            out.ldy = (long long) (((unsigned long long) d0.w << 32) | d0.z);
            out.tag = (unsigned long long *) (((unsigned long long) d2.y << 32) | d2.x);
            out.tag_stride = a.tag_stride; out.epoch = epoch0 + (unsigned) n + 1u;
            out.world = a.world; out.delta = a.delta;
            const int type = (int) d1.x, nrows = (int) d1.y, rot = (int) d1.w;
            const int nunits = nrows * upr;

            if (a.debug & 2) {
                __syncwarp();
                if (segs > 1) { if (lane == 0) mbar_arrive(part_bar(st)); pmask ^= 1u << st; }
                if (lane == 0) mbar_arrive(empty_bar(st));
            } else if (segs == 1) {
                int t0 = warp - rot; if (t0 < 0) t0 += NW;
                for (int row0 = t0 * rp; row0 < nrows; row0 += NW * rp) {
                    const int row = row0 + sub;
                    const bool live = active && row < nrows;
                    float v = 0.0f;
                    if (live) v = type == B200_TYPE_Q4_K ? unit_q4k_r(stage + (uint32_t) (row * upr + b) * 144u, x) : unit_q6k_r(stage, nunits, row * upr + b, x);
                    for (int o = 1; o < upr; o <<= 1) {              // segmented: lanes of one row are consecutive, fixed order
                        const float w = __shfl_down_sync(0xffffffffu, v, o);
                        if (b + o < upr) v += w;
                    }
                    if (live && b == 0) put_y(out, 0, row, v);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(empty_bar(st));
            } else {
                const uint32_t part = part0 + (uint32_t) (st * CX_PART_FLOATS * 4);
                int r0 = idx - rot; if (r0 < 0) r0 += per;
                for (int r = r0; r < nrows; r += per) {
                    float v = 0.0f;
                    if (active) v = type == B200_TYPE_Q4_K ? unit_q4k_r(stage + (uint32_t) (r * upr + b) * 144u, x) : unit_q6k_r(stage, nunits, r * upr + b, x);
                    v = warp_sum(v);
                    if (lane == 0) stsf(part + (uint32_t) (r * segs + seg) * 4, v);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(part_bar(st));            // release: the partial sums above are visible to whoever sees the phase
                if (seg == segs - 1) {                               // the last class adds the pieces of every row in a fixed order
                    mbar_wait(part_bar(st), (pmask >> st) & 1u);
                    if (!(a.debug & 8)) {
                        for (int r = idx + per * lane; r < nrows; r += per * 32) {
                            float v = 0.0f;
                            for (int sg = 0; sg < segs; ++sg) v += ldsf(part + (uint32_t) (r * segs + sg) * 4);
                            put_y(out, 0, r, v);
                        }
                    }
                    __syncwarp();
                }
                pmask ^= 1u << st;
                if (lane == 0) mbar_arrive(empty_bar(st));
            }
            if (++st == NS) { st = 0; ph ^= 1u; }
        }
        if (tl && tid == 0) { tl[n * 8 + 4] = c_wait; tl[n * 8 + 5] = clock64() - c_loop; }
        if (nd.signal) {
            // the next node waits on the arrival counter: every store of this CTA is issued before the arrival (a release publishes them)
            const long long c_rel = tl ? clock64() : 0;
            cons_bar_x();
            if (tid == 0) {
                if (nd.signal == 3) asm volatile("red.relaxed.gpu.global.add.u64 [%0], 1;" :: "l"(a.counter_local) : "memory");
                else if (a.world > 1) {
                    if (nd.signal == 2) asm volatile("fence.acq_rel.sys;" ::: "memory");
                    for (int g = 0; g < a.world; ++g)
                        asm volatile("red.relaxed.sys.global.add.u64 [%0], 1;" :: "l"(reinterpret_cast<char *>(a.counter) + a.delta[g]) : "memory");
                } else if (nd.signal == 2) asm volatile("red.release.gpu.global.add.u64 [%0], 1;" :: "l"(a.counter) : "memory");
                else                       asm volatile("red.relaxed.gpu.global.add.u64 [%0], 1;" :: "l"(a.counter) : "memory");
                if (tl) tl[n * 8 + 6] = clock64() - c_rel;
            }
        }
        if (tl && tid == 0) { tl[n * 8 + 3] = clock64(); unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); tl[n * 8 + 7] = (long long) t; }
        nd = nd_next; cnt = cnt_next;
    }
    if (a.world > 1) {
        cons_bar_x();
        if (tid == 0) {
            asm volatile("fence.acq_rel.sys;" ::: "memory");
            for (int g = 0; g < a.world; ++g)
                asm volatile("red.relaxed.sys.global.add.u64 [%0], 1;" :: "l"(reinterpret_cast<char *>(a.done) + a.delta[g]) : "memory");
            if (blockIdx.x == 0) {
                const unsigned long long want = (launch + 1ull) * (unsigned long long) gridDim.x * (unsigned) a.world;
                unsigned long long v, t0 = 0;
                for (unsigned spin = 0;; ++spin) {
                    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(a.done) : "memory");
                    if (v >= want) break;
                    __nanosleep(100);
                    if ((spin & 1023u) == 1023u) {
                        unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                        if (t0 == 0) t0 = t; else if (t - t0 > 4000000000ull) __trap();
                    }
                }
            }
        }
    }
    if (blockIdx.x == 0 && tid == 0) *a.launches = launch + 1ull;
}
