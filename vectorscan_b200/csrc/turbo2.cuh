// turbo2.cuh -- k_turbo2, the literal front end (included by hs_gpu.cu after turbo.cuh, whose bulk-copy helpers,
// second-tier table, slot-table confirm and ordered drain it reuses).  Tables: turbo_tables.h ("gram table").
//
// What changed against k_turbo, and why (profiles/r01_ncu_k_turbo_*: 12.6 / 19.9 thread-instructions per byte, 55-58 %
// of the shared-memory wavefronts were bank-conflict replays of random 8-byte reads):
//   * first tier = ONE 4-byte shared-memory read per two text bytes (a random LDS.32 costs ~3.5 wavefronts per warp, a
//     random LDS.64 ~6.4): stride 1 tests two end positions against the union set of their shared row; stride 2 (literal
//     sets without many shortest-possible literals) looks at every other position only;
//   * the row hash is a multiply of the unmasked text window (turbo_tables.h), the shift-amount registers are the
//     windows themselves (a funnel shift looks at 5 bits): ~8 instructions per lookup, no PRMT packing, no masks;
//   * nothing after the first tier runs lane-local any more.  Survivors are compacted, in position order, into a
//     per-warp ring of bare positions; whenever the ring holds 32, ONE full-warp pass (a lane per candidate) fetches the
//     8-byte window from global memory (L2: the step just came through it), applies the tail table and queues what
//     is left for the slot-table confirm, again 32 at a time.  Rare work therefore costs instructions in proportion to
//     the candidates, not one divergent trip per step;
//   * the text registers are loaded at the top of a step and the step's buffer is handed back to the bulk-copy engine at
//     once (the later tiers do not read it), so every warp keeps almost two steps of HBM latency covered.
#pragma once

constexpr int T2Q1 = 128;                  // ring of first-tier survivors (positions), power of two
constexpr int T2E = 1 + TCHUNK / 4;        // extended chunk words: the 4 bytes before the chunk, then the chunk

template <bool ORDERED>
struct Turbo2WarpShared {
    __align__(16) u8 text[TNBUF][TBUF_BYTES];
    unsigned long long mbar[TNBUF];
    u32 q1[T2Q1];
    uint2 q2[TQ2CAP];
    Rec stage[ORDERED ? 1 : TSTAGE];
    u32 stage_n;
    u32 pad[3];
};
template <bool ORDERED> __host__ __device__ constexpr int turbo2_max_warps() { return 32; }

template <bool ORDERED>
__host__ __device__ inline size_t turbo2_smem_bytes(const DevDb &db, int warps) {
    return ((size_t)4 << db.gG1) + ((size_t)4 << (db.tBL - 5)) + sizeof(Turbo2WarpShared<ORDERED>) * (size_t)warps;
}

// 32-bit text window starting at chunk byte `o` (o >= -4): its low byte is text[o]
template <int o>
__device__ __forceinline__ u32 t2_win(const u32 (&E)[T2E]) {
    constexpr int e = o + 4, w = e >> 2;
    static_assert(e >= 0 && w < T2E, "window outside the extended chunk");
    if ((e & 3) == 0) return E[w];
    return __funnelshift_r(E[w], w + 1 < T2E ? E[w + 1] : 0u, 8 * (e & 3));
}

// stride 1: pairs of end positions (i-1, i), i = 1, 3, ..., 15; the candidate mask is built first-position-on-top
template <int K, int i>
struct Turbo2Pairs {
    static __device__ __forceinline__ void run(const u32 *tab, u32 mul, u32 rshift, const u32 (&E)[T2E], u32 &acc) {
        const u32 ent = K > 1 ? tab[(t2_win<i - K + 1>(E) * mul) >> rshift] : tab[0];
        // sets are stored bit-reversed: shifting left by low5(byte) brings the byte's flag to bit 31, a funnel shift by
        // one appends it to the mask
        const u32 ra = __funnelshift_l(0u, ent, t2_win<i - K>(E));   // position i-1: byte before the shared ones
        acc = __funnelshift_l(ra, acc, 1);
        const u32 rb = __funnelshift_l(0u, ent, t2_win<i>(E));       // position i: byte after them
        acc = __funnelshift_l(rb, acc, 1);
        Turbo2Pairs<K, i + 2>::run(tab, mul, rshift, E, acc);
    }
};
template <int K>
struct Turbo2Pairs<K, TCHUNK + 1> {
    static __device__ __forceinline__ void run(const u32 *, u32, u32, const u32 (&)[T2E], u32 &) {}
};

// stride 2: grams ending at the odd chunk positions i = 1, 3, ..., 15; one flag per sample, first sample on top
template <int K, int i>
struct Turbo2Samples {
    static __device__ __forceinline__ void run(const u32 *tab, u32 mul, u32 rshift, const u32 (&E)[T2E], u32 &acc) {
        const u32 ent = tab[(t2_win<i - K + 1>(E) * mul) >> rshift];
        const u32 r = __funnelshift_l(0u, ent, t2_win<i>(E));
        acc = __funnelshift_l(r, acc, 1);
        Turbo2Samples<K, i + 2>::run(tab, mul, rshift, E, acc);
    }
};
template <int K> struct Turbo2Samples<K, TCHUNK + 1> { static __device__ __forceinline__ void run(const u32 *, u32, u32, const u32 (&)[T2E], u32 &) {} };

// first tier of one chunk.  S == 1: 16 flags, bit 15 - j = end position j.  S == 2: 8 flags, bit 7 - k = sample k
// (chunk position 2k + 1).
template <int K, int S>
__device__ __forceinline__ u32 turbo2_chunk(const u32 *tab, u32 mul, u32 rshift, const u32 (&E)[T2E]) {
    u32 acc = 0;
    if (S == 1) Turbo2Pairs<K, 1>::run(tab, mul, rshift, E, acc);
    else Turbo2Samples<K, 1>::run(tab, mul, rshift, E, acc);
    return acc;
}

// Second tier + hand-over to the confirm queue for up to 32 ring entries (a lane each), warp-synchronous.
// Ring entry = end position relative to qbase (= span start - THALO, so that it is never negative).  Survivors go to q2
// in position order as {position, K-byte tail}; q2 is confirmed 32 at a time.
template <int K, bool ORDERED>
__device__ __forceinline__ void turbo2_tier2(const Ctx &c, const u32 *tail_tab, const u32 *q1, u32 head, u32 count, uint2 *q2,
                                             u32 &q2n, u64 qbase, u32 &n_probe, Rec *seg, u32 &cursor) {
    const u32 ln = lane_id(), word_shift = 37 - c.db.tBL;
    const bool has_long = c.db.tL != 0;
    bool keep = false;
    u32 pos = 0, ktail = 0;
    if (ln < count) {
        pos = q1[(head + ln) & (T2Q1 - 1)];
        const u64 w = load_tail8(c.cp.text, qbase + pos);
        ktail = (u32)(w >> (64 - 8 * K));
        keep = turbo_tail_hit(tail_tab, word_shift, has_long, ktail, (u32)(w >> (56 - 8 * K)));
    }
    const u32 b = __ballot_sync(0xffffffffu, keep);
    if (!b) return;
    if (keep) q2[q2n + __popc(b & ((1u << ln) - 1))] = make_uint2(pos, ktail);
    q2n += __popc(b);
    __syncwarp();
    if (q2n >= 32) {  // fewer than 32 on entry, at most 32 added
        turbo_drain<ORDERED>(c, q2, qbase, 32, n_probe, seg, cursor);
        const u32 rem = q2n - 32;
        const uint2 k0 = ln < rem ? q2[32 + ln] : make_uint2(0, 0);
        __syncwarp();
        if (ln < rem) q2[ln] = k0;
        __syncwarp();
        q2n = rem;
    }
}

template <int K, int S, bool ORDERED>
__global__ void __launch_bounds__(1024, 1) k_turbo2(const DevDb db, const DevCorpus cp, const DevArena ar) {
    using WS = Turbo2WarpShared<ORDERED>;
    extern __shared__ __align__(16) u8 smem[];
    u32 *tab = reinterpret_cast<u32 *>(smem);
    const u32 rows = 1u << db.gG1, bloom_words = 1u << (db.tBL - 5);
    u32 *bloom = tab + rows;
    WS *wsh = reinterpret_cast<WS *>(smem + (size_t)rows * 4 + (size_t)bloom_words * 4);
    const u32 TW = blockDim.x >> 5, warp = threadIdx.x >> 5, ln = lane_id();
    WS &ws = wsh[warp];
    // every warp of the grid takes one contiguous span of 1 KiB steps, balanced to within one step
    const u64 steps_total = (cp.total + TSTEP_BYTES - 1) / TSTEP_BYTES;
    const u64 W = (u64)gridDim.x * TW, w = (u64)blockIdx.x * TW + warp;
    const u64 r0 = steps_total * w / W, r1 = steps_total * (w + 1) / W;
    const u64 span_base = r0 * TSTEP_BYTES;
    const u64 qbase = span_base - THALO;           // queue positions are relative to this (wraps below zero for the first span)
    const u32 nsteps = (u32)(r1 - r0);
    const u8 *gsrc = cp.text + span_base - THALO;  // FRONT_PAD keeps the first halo in bounds, BACK_PAD the last step
    if (ln == 0) {
#pragma unroll
        for (int b = 0; b < TNBUF; b++) mbar_init(&ws.mbar[b], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#pragma unroll
        for (int b = 0; b < TNBUF; b++)
            if ((u32)b < nsteps) bulk_load(ws.text[b], gsrc + (size_t)b * TSTEP_BYTES, TBUF_BYTES, &ws.mbar[b]);
        ws.stage_n = 0;
    }
    {
        // every CTA copies the same tables: each starts at a different offset so that the SMs do not all ask the same
        // L2 slice for the same line at the same moment
        const uint4 *src = reinterpret_cast<const uint4 *>(db.gram_tab);
        uint4 *dst = reinterpret_cast<uint4 *>(tab);
        const u32 n1 = rows / 4, rot1 = (u32)(((u64)blockIdx.x * n1) / gridDim.x) & ~31u;
#pragma unroll 4
        for (u32 i = threadIdx.x; i < n1; i += blockDim.x) {
            const u32 j = (i + rot1) & (n1 - 1);
            dst[j] = __ldg(src + j);
        }
        const uint4 *bsrc = reinterpret_cast<const uint4 *>(db.bloom);
        uint4 *bdst = reinterpret_cast<uint4 *>(bloom);
        const u32 n2 = bloom_words / 4, rot2 = (u32)(((u64)blockIdx.x * n2) / gridDim.x) & ~31u;
#pragma unroll 4
        for (u32 i = threadIdx.x; i < n2; i += blockDim.x) {
            const u32 j = (i + rot2) & (n2 - 1);
            bdst[j] = __ldg(bsrc + j);
        }
    }
    __syncthreads();

    const Ctx c{db, cp, ar, nullptr, ws.stage, &ws.stage_n, ORDERED ? 0u : (u32)TSTAGE, nullptr, nullptr};
    u32 *q1 = ws.q1;
    uint2 *q2 = ws.q2;
    const u32 mul = db.gmul, rshift = 32 - db.gG1;
    u32 n_cand = 0, n_probe = 0;
    Rec *seg = ORDERED ? ar.recs + (size_t)w * ar.seg_cap : nullptr;
    u32 cursor = 0, q1h = 0, q1n = 0, q2n = 0;
    u32 s2_carry = 1;  // sampled form: "the sample just before this chunk hit"; unknown at the span's start, so assume it
    const u32 lt = (1u << ln) - 1;
#pragma unroll 1
    for (u32 step = 0; step < nsteps; step++) {
        const u32 b = step % TNBUF;
        mbar_wait(&ws.mbar[b], (step / TNBUF) & 1);
        u32 m;
        {
            const u8 *cb = ws.text[b] + THALO + ln * TCHUNK;
            const uint4 c0 = *reinterpret_cast<const uint4 *>(cb), c1 = *reinterpret_cast<const uint4 *>(cb + TROW_BYTES);
            const u32 p0 = *reinterpret_cast<const u32 *>(cb - 4), p1 = *reinterpret_cast<const u32 *>(cb + TROW_BYTES - 4);
            __syncwarp();
            // the registers hold the step: its buffer can take the step TNBUF ahead already
            if (ln == 0 && step + TNBUF < nsteps)
                bulk_load(ws.text[b], gsrc + (size_t)(step + TNBUF) * TSTEP_BYTES, TBUF_BYTES, &ws.mbar[b]);
            const u32 E0[T2E] = {p0, c0.x, c0.y, c0.z, c0.w};
            m = turbo2_chunk<K, S>(tab, mul, rshift, E0);
            const u32 E1[T2E] = {p1, c1.x, c1.y, c1.z, c1.w};
            m = (m << 16) | turbo2_chunk<K, S>(tab, mul, rshift, E1);
        }
        const u32 step_pos = step * TSTEP_BYTES;
        const bool last_step = span_base + step_pos + TSTEP_BYTES > cp.total;  // the corpus ends inside this step
        const long long left0 = (long long)cp.total - (long long)(span_base + step_pos + ln * TCHUNK), left1 = left0 - TROW_BYTES;
        if (S == 2) {
            // m: bit 23 - k / bit 7 - k = sample k (chunk position 2k + 1) of row 0 / row 1.  A hit at p means "a
            // literal may end at p - 1, p or p + 1": spread the 8 flags of each row over the odd positions, then let
            // every flag also mark its two neighbours.  The flag that reaches over the chunk's end arrives at the next
            // chunk (next lane, next row, next step) as `pin`.
            if (last_step) {  // keep the samples up to and INCLUDING the position one past the corpus
                const long long s0 = left0 < 1 ? 0 : (left0 + 1) / 2, s1 = left1 < 1 ? 0 : (left1 + 1) / 2;
                const u32 k0 = s0 >= 8 ? 0xffu : (0xffu << (8 - (int)s0)) & 0xffu, k1 = s1 >= 8 ? 0xffu : (0xffu << (8 - (int)s1)) & 0xffu;
                m &= (k0 << 16) | k1;
            }
            const u32 mu = __shfl_up_sync(0xffffffffu, m, 1), m31 = __shfl_sync(0xffffffffu, m, 31);
            const u32 pin0 = ln ? (mu >> 16) & 1u : s2_carry, pin1 = ln ? mu & 1u : (m31 >> 16) & 1u;
            s2_carry = m31 & 1u;
            u32 x = m & 0x00ff00ffu;
            x = (x | (x << 4)) & 0x0f0f0f0fu;
            x = (x | (x << 2)) & 0x33333333u;
            x = (x | (x << 1)) & 0x55555555u;   // sample k of row 0 at bit 30 - 2k, of row 1 at bit 14 - 2k
            m = x | (x << 1) | ((x >> 1) & 0x7fff7fffu) | (pin0 << 31) | (pin1 << 15);
        }
        // m: bit 31 - j = position j of the lane's row-0 chunk, bit 15 - j = position j of its row-1 chunk
        if (last_step) {  // clip end positions beyond the corpus
            const u32 k0 = left0 >= 16 ? 0xffffu : left0 <= 0 ? 0u : (0xffffu << (16 - (int)left0)) & 0xffffu;
            const u32 k1 = left1 >= 16 ? 0xffffu : left1 <= 0 ? 0u : (0xffffu << (16 - (int)left1)) & 0xffffu;
            m &= (k0 << 16) | k1;
        }
        if (!__any_sync(0xffffffffu, m != 0)) continue;
        n_cand += __popc(m);
        // ---- compaction into the ring, position order: row 0 lane-major, then row 1 ----
        const u32 n0 = __popc(m >> 16), n1 = __popc(m & 0xffffu);
        u32 slot0, slot1, total;
        if (!__any_sync(0xffffffffu, (n0 | n1) > 1)) {
            const u32 b0 = __ballot_sync(0xffffffffu, n0 != 0), b1 = __ballot_sync(0xffffffffu, n1 != 0);
            slot0 = __popc(b0 & lt);
            slot1 = __popc(b0) + __popc(b1 & lt);
            total = __popc(b0) + __popc(b1);
        } else {
            const u32 n01 = n0 | (n1 << 16);
            u32 incl = n01;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const u32 t = __shfl_up_sync(0xffffffffu, incl, d);
                if (ln >= (u32)d) incl += t;
            }
            const u32 tot = __shfl_sync(0xffffffffu, incl, 31);
            slot0 = (incl & 0xffffu) - n0;
            slot1 = (tot & 0xffffu) + (incl >> 16) - n1;
            total = (tot & 0xffffu) + (tot >> 16);
        }
        u32 done = 0;  // survivors of this step already moved to the ring (warp-uniform)
        while (done < total) {
            const u32 room = T2Q1 - q1n, lim = done + room;  // this pass takes ranks [done, lim)
            u32 mm = m, s0 = slot0, s1 = slot1;
            while (mm) {
                const u32 lz = __clz(mm);
                mm &= ~(0x80000000u >> lz);
                const bool second = lz >= 16;
                const u32 rank = second ? s1++ : s0++;
                if (rank >= done && rank < lim) {
                    const u32 pos = ln * TCHUNK + (lz & 15) + (second ? TROW_BYTES : 0);
                    q1[(q1h + q1n + rank - done) & (T2Q1 - 1)] = step_pos + pos + THALO;
                }
            }
            const u32 took = total - done < room ? total - done : room;
            q1n += took;
            done += took;
            __syncwarp();
            while (q1n >= 32) {
                turbo2_tier2<K, ORDERED>(c, bloom, q1, q1h, 32, q2, q2n, qbase, n_probe, seg, cursor);
                q1h = (q1h + 32) & (T2Q1 - 1);
                q1n -= 32;
            }
        }
    }
    if (q1n) turbo2_tier2<K, ORDERED>(c, bloom, q1, q1h, q1n, q2, q2n, qbase, n_probe, seg, cursor);
    if (q2n) turbo_drain<ORDERED>(c, q2, qbase, q2n, n_probe, seg, cursor);
    if (ORDERED) {
        if (ln == 0) ar.seg_count[w] = cursor;
    }
    if (!ORDERED) flush_stage(c);
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        n_probe += __shfl_down_sync(0xffffffffu, n_probe, d);
        n_cand += __shfl_down_sync(0xffffffffu, n_cand, d);
    }
    if (ln == 0) {
        if (n_cand) atomicAdd(&ar.stats[0], (unsigned long long)n_cand);
        if (n_probe) atomicAdd(&ar.stats[1], (unsigned long long)n_probe);
    }
}
