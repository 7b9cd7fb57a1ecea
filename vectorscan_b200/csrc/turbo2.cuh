// turbo2.cuh -- k_turbo2, the literal front end (included by hs_gpu.cu after turbo.cuh, whose bulk-copy helpers,
// second-tier table, slot-table confirm and ordered drain it reuses).  Tables: turbo_tables.h ("gram table").
//
// What changed against k_turbo, and why (profiles/r01_ncu_k_turbo_*: 12.6 / 19.9 thread-instructions per byte, 55-58 %
// of the shared-memory wavefronts were bank-conflict replays of random 8-byte reads):
//   * first tier = ONE 4-byte shared-memory read per two text bytes (a random LDS.32 costs ~3.5 wavefronts per warp, a
//     random LDS.64 ~6.4): two end positions are tested against the union set of their shared row.  Sets whose union
//     form would pass too many positions stay with k_turbo's 8-byte rows (turbo_tables.h, `wide`).  A form that looks at
//     every other position only was measured and dropped: it saves two shifts per lookup and triples the second tier's
//     work;
//   * the row hash is a multiply of the unmasked text window (turbo_tables.h), the shift-amount registers are the
//     windows themselves (a funnel shift looks at 5 bits): ~8 instructions per lookup, no PRMT packing, no masks;
//   * nothing after the first tier runs lane-local any more.  Survivors are compacted, in position order, into a
//     per-warp queue of bare positions; whenever it holds 32, ONE full-warp pass (a lane per candidate) fetches the
//     K + 1 bytes ending there from global memory (L2: the step just came through it), applies the tail table and queues what
//     is left for the slot-table confirm, again 32 at a time.  Rare work therefore costs instructions in proportion to
//     the candidates, not one divergent trip per step;
//   * the text registers are loaded at the top of a step and the step's buffer is handed back to the bulk-copy engine at
//     once (the later tiers do not read it), so every warp keeps almost two steps of HBM latency covered.
#pragma once

constexpr int T2Q1 = 128;                  // first-tier survivors (positions) waiting for the second tier
constexpr int T2LANE = 2 * TCHUNK;         // every lane owns 32 consecutive bytes of a step
constexpr int T2MAXW = 24;                 // warps per CTA at most (768 threads: 85 registers each)
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

// pairs of end positions (i-1, i), i = 1, 3, ..., 15; the candidate mask is built first-position-on-top
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

// first tier of one chunk: 16 flags, bit 15 - j = end position j
template <int K>
__device__ __forceinline__ u32 turbo2_chunk(const u32 *tab, u32 mul, u32 rshift, const u32 (&E)[T2E]) {
    u32 acc = 0;
    Turbo2Pairs<K, 1>::run(tab, mul, rshift, E, acc);
    return acc;
}

// Second tier + hand-over to the confirm queue for up to 32 entries q1[head ...) (a lane each), warp-synchronous.
// Queue entry = end position relative to qbase (= span start - THALO, so that it is never negative).  Survivors go to q2
// in position order as {position, K-byte tail}; q2 is confirmed 32 at a time.  (Issuing a batch's window loads one batch
// ahead of their use was measured: the extra live registers cost more than the L2 round trip they hid.)
template <int K, bool ORDERED>
__device__ __forceinline__ void turbo2_tier2(const Ctx &c, const u32 *tail_tab, const u32 *q1, u32 head, u32 count, uint2 *q2,
                                             u32 &q2n, u64 qbase, u32 &n_probe, Rec *seg, u32 &cursor) {
    const u32 ln = lane_id(), word_shift = 37 - c.db.tBL;
    const bool has_long = c.db.tL != 0;
    bool keep = false;
    u32 pos = 0, ktail = 0;
    if (ln < count) {
        pos = q1[head + ln];
        // the K + 1 bytes ending at the position lie inside two aligned words (L2: the step just came through it)
        const u8 *first = c.cp.text + (qbase + pos) - K;
        const u32 o = (u32)(reinterpret_cast<uintptr_t>(first) & 3);
        const u32 *wp = reinterpret_cast<const u32 *>(first - o);
        const u32 w0 = __ldg(wp), w1 = __ldg(wp + 1);
        const u32 extra = __funnelshift_r(w0, w1, 8 * o);   // low byte = the byte before the K-tail
        if (K == 4) ktail = o == 3 ? w1 : __funnelshift_r(w0, w1, 8 * o + 8);
        else ktail = (extra >> 8) & ((1u << (8 * K)) - 1);
        keep = turbo_tail_hit(tail_tab, word_shift, has_long, ktail, extra);
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

template <int K, bool ORDERED>
__global__ void __launch_bounds__(T2MAXW * 32, 1) k_turbo2(const DevDb db, const DevCorpus cp, const DevArena ar) {
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
    u32 cursor = 0, q1n = 0, q2n = 0;
    const u32 lt = (1u << ln) - 1;
    const u64 corpus_left = cp.total - span_base;   // text bytes from the span's start to the corpus end
#pragma unroll 1
    for (u32 step = 0; step < nsteps; step++) {
        const u32 b = step % TNBUF;
        mbar_wait(&ws.mbar[b], (step / TNBUF) & 1);
        // the lane's 32 consecutive bytes, and the 4 before them
        u32 m;
        {
            const u8 *cb = ws.text[b] + THALO + ln * T2LANE;
            const uint4 c0 = *reinterpret_cast<const uint4 *>(cb), c1 = *reinterpret_cast<const uint4 *>(cb + TCHUNK);
            // the 4 bytes before the lane's: the previous lane's last word; lane 0 reads them from the halo
            u32 p0 = __shfl_up_sync(0xffffffffu, c1.w, 1);
            if (ln == 0) p0 = *reinterpret_cast<const u32 *>(cb - 4);
            __syncwarp();
            // the registers hold the step: its buffer can take the step TNBUF ahead already
            if (ln == 0 && step + TNBUF < nsteps)
                bulk_load(ws.text[b], gsrc + (size_t)(step + TNBUF) * TSTEP_BYTES, TBUF_BYTES, &ws.mbar[b]);
            const u32 E0[T2E] = {p0, c0.x, c0.y, c0.z, c0.w};
            m = turbo2_chunk<K>(tab, mul, rshift, E0);
            const u32 E1[T2E] = {c0.w, c1.x, c1.y, c1.z, c1.w};
            m = (m << 16) | turbo2_chunk<K>(tab, mul, rshift, E1);
        }
        const u32 step_pos = step * TSTEP_BYTES;
        const bool last_step = (u64)step_pos + TSTEP_BYTES > corpus_left;  // the corpus ends inside this step
        const long long left = (long long)corpus_left - (long long)(step_pos + ln * T2LANE);  // lane's bytes inside the corpus
        // m: bit 31 - j = end position j of the lane's 32 bytes
        if (last_step) m &= left >= 32 ? 0xffffffffu : left <= 0 ? 0u : 0xffffffffu << (32 - (int)left);
        if (!__any_sync(0xffffffffu, m != 0)) continue;
        // ---- compaction into q1, position order = lane order ----
        const u32 n = __popc(m);
        n_cand += n;
        u32 excl, total;
        if (!__any_sync(0xffffffffu, n >= 8)) {
            // the usual case: three independent ballots (one per bit of n) instead of five dependent shuffles
            const u32 b0 = __ballot_sync(0xffffffffu, n & 1u), b1 = __ballot_sync(0xffffffffu, n & 2u), b2 = __ballot_sync(0xffffffffu, n & 4u);
            excl = __popc(b0 & lt) + 2 * __popc(b1 & lt) + 4 * __popc(b2 & lt);
            total = __popc(b0) + 2 * __popc(b1) + 4 * __popc(b2);
        } else {
            u32 incl = n;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const u32 t = __shfl_up_sync(0xffffffffu, incl, d);
                if (ln >= (u32)d) incl += t;
            }
            total = __shfl_sync(0xffffffffu, incl, 31);
            excl = incl - n;
        }
        const u32 pos_top = step_pos + ln * T2LANE + THALO + 31;  // queue position of the lane's byte j is pos_top - (31 - j)
        if (q1n + total <= T2Q1) {
            u32 *pf = q1 + q1n + excl;
            u32 mm = m;
            while (mm) {
                const u32 f = 31 - __clz(mm);  // bit index of the first remaining position
                *pf++ = pos_top - f;
                mm &= ~(1u << f);
            }
            q1n += total;
            __syncwarp();
            u32 at = 0;
            for (; at + 32 <= q1n; at += 32) turbo2_tier2<K, ORDERED>(c, bloom, q1, at, 32, q2, q2n, qbase, n_probe, seg, cursor);
            if (at) {  // what is left (fewer than 32) moves to the front
                const u32 rem = q1n - at, v = ln < rem ? q1[at + ln] : 0u;
                __syncwarp();
                if (ln < rem) q1[ln] = v;
                q1n = rem;
                __syncwarp();
            }
        } else {
            // a step with more survivors than the queue holds (dense matches): empty the queue, then feed the step
            // through it 96 ranks at a time
            if (q1n) turbo2_tier2<K, ORDERED>(c, bloom, q1, 0, q1n, q2, q2n, qbase, n_probe, seg, cursor);
            q1n = 0;
            for (u32 lo = 0; lo < total; lo += 96) {
                u32 mm = m, r = excl;
                while (mm) {
                    const u32 f = 31 - __clz(mm);
                    if (r >= lo && r < lo + 96) q1[r - lo] = pos_top - f;
                    r++;
                    mm &= ~(1u << f);
                }
                __syncwarp();
                const u32 have = total - lo < 96 ? total - lo : 96;
                for (u32 at = 0; at < have; at += 32)
                    turbo2_tier2<K, ORDERED>(c, bloom, q1, at, have - at < 32 ? have - at : 32, q2, q2n, qbase, n_probe, seg, cursor);
                __syncwarp();
            }
        }
    }
    if (q1n) turbo2_tier2<K, ORDERED>(c, bloom, q1, 0, q1n, q2, q2n, qbase, n_probe, seg, cursor);
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
