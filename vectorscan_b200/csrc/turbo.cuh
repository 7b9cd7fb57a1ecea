// turbo.cuh -- the "turbo" front end and its confirm (included by hs_gpu.cu; see turbo_tables.h for the
// tables).  Same strip/row/chunk geometry as k_scan; differences:
//   * one 8-byte shared-memory lookup filters TWO end positions (pair table), ~5 ALU ops per position;
//   * 24 warps per SM share the one table, the next row's LDG.128 is issued before the current row is
//     processed (software prefetch: ~24 x 2 x 512 B in flight per SM);
//   * tier-1 survivors are queued as bare positions; in batches of 32 (one lane each, all lanes busy) they go
//     through the tail bloom (tier 2, shared memory); its survivors, again in batches of 32, reload the 8-byte
//     window from L1/L2, probe the slot table (one 32-byte L2 read pair per probe) and compare
//     (window & msk) == v exactly as confWithBit does; only then is the buffer looked up for the size check.
#pragma once

constexpr int TQ2CAP = 64;   // tier-2 survivors awaiting the slot-table confirm (drained 32 at a time)
constexpr int TSTAGE = 32;   // staged match records per warp (general mode)
constexpr int TCHUNK = 16;                 // bytes per lane per row (one LDS.128)
constexpr int TROW_BYTES = 32 * TCHUNK;    // 512
constexpr int TSTEP_BYTES = 2 * TROW_BYTES;  // a warp step = two rows = 1 KiB, fetched by one bulk copy
constexpr int THALO = 16;                  // bytes before the step fetched with it (look-behind of tier 1 / tier 2)
constexpr int TBUF_BYTES = THALO + TSTEP_BYTES;
constexpr int TNBUF = 2;                   // step buffers per warp (one being scanned, one in flight)
constexpr int TEW = 2 + TCHUNK / 4;        // extended chunk words: 8 bytes before the chunk, then the chunk

// Per-warp shared memory.  ORDERED mode needs no record staging, so it affords 24 warps per SM; the general
// mode runs 20.
template <bool ORDERED>
struct TurboWarpShared {
    __align__(16) u8 text[TNBUF][TBUF_BYTES];  // [THALO bytes before the step][the step], filled by cp.async.bulk
    unsigned long long mbar[TNBUF];            // one mbarrier per buffer (transaction-count completion)
    uint2 q2[TQ2CAP];                          // {span-relative position, K-byte tail}
    Rec stage[ORDERED ? 1 : TSTAGE];
    u32 stage_n;
    u32 pad[3];
};
template <bool ORDERED> __host__ __device__ constexpr int turbo_warps() { return ORDERED ? 24 : 20; }

template <bool ORDERED>
__host__ __device__ inline size_t turbo_smem_bytes(const DevDb &db) {
    return ((size_t)8 << db.tH1) + ((size_t)4 << (db.tBL - 5)) + sizeof(TurboWarpShared<ORDERED>) * turbo_warps<ORDERED>();
}

// ---- bulk asynchronous copy (TMA, non-tensor form) + mbarrier, sm_90+/sm_100a PTX -------------------------
__device__ __forceinline__ u32 smem_u32(const void *p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// one thread: expect `bytes` on the barrier and start the copy global -> shared that will deliver them
__device__ __forceinline__ void bulk_load(void *dst, const void *src, u32 bytes, unsigned long long *bar) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // earlier generic-proxy accesses to dst are done
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, u32 parity) {
    u32 done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!done);
}

__device__ __forceinline__ void turbo_hit(const Ctx &c, const uint4 &b, u64 g) {
    const BufRef br = find_buffer(c.cp, 0, g);
    if (!passes_gates(c.db, br.len)) return;
    const u32 i = (u32)(g - br.start);
    const u32 size = b.w & 0xff, kind = (b.w >> 8) & 0xf;
    if (size > i + 1) return;  // the literal would start before its buffer (fdr_confirm_runtime.h:77-88)
    const u8 *btext = c.cp.text + br.start;
    if (kind == turbo::SLOT_REPORT) {
        emit(c, br.idx, i + 1, b.y, 0);
    } else if (kind == turbo::SLOT_PROGRAM) {
        run_program(c, b.y, br.idx, btext, br.len, i + 1);
    } else if (check_literal(c.db.bc, b.z, b.w >> 16, kind == turbo::SLOT_LIT_NOCASE_REPORT, btext, i + 1)) {
        emit(c, br.idx, i + 1, b.y, 0);
    }
}

// One tier-2 survivor: global end position g, key = its K-byte tail (so the slot probe does not wait for the
// window).  Slots h and h + 1 are fetched together: with the table a quarter full most chains end there.
__device__ __forceinline__ void turbo_confirm(const Ctx &c, u64 g, u32 key, u32 &n_probe) {
    const u32 mask = (1u << c.db.tH2) - 1;
    u32 h = turbo::slot_home(key, c.db.tH2);
    const uint4 *slots = reinterpret_cast<const uint4 *>(c.db.slots);
    uint4 a0 = __ldg(slots + 2 * (size_t)h), b0 = __ldg(slots + 2 * (size_t)h + 1);
    uint4 a1 = __ldg(slots + 2 * (size_t)((h + 1) & mask)), b1 = __ldg(slots + 2 * (size_t)((h + 1) & mask) + 1);
    const u64 w = load_tail8(c.cp.text, g);
    for (;;) {
        // slot = (v, msk | key, target, lit_off, meta)
        if (!((b0.w >> 8) & 0xf)) break;
        n_probe++;
        if (b0.x == key) {
            const u64 v = ((u64)a0.y << 32) | a0.x, m = ((u64)a0.w << 32) | a0.z;
            if ((w & m) == v) turbo_hit(c, b0, g);
        }
        h = (h + 1) & mask;
        a0 = a1; b0 = b1;
        if (!((b0.w >> 8) & 0xf)) break;
        a1 = __ldg(slots + 2 * (size_t)((h + 1) & mask));
        b1 = __ldg(slots + 2 * (size_t)((h + 1) & mask) + 1);
    }
}

// Second tier: the tail table (turbo_tables.h, 3.).  One shared-memory word per position: true when some
// literal's (K+1)-byte tail -- or K-byte tail, for a literal that short -- may end at the position.
// ktail = the position's K-byte tail (first byte lowest); extra = a register whose low byte is the byte before it.
__device__ __forceinline__ bool turbo_tail_hit(const u32 *tail_tab, u32 word_shift, bool has_long, u32 ktail, u32 extra) {
    const u32 h = turbo::tail_hash(ktail);
    const u32 w = tail_tab[h >> word_shift];
    // funnel shifts look at the low 5 bits of the shift register only (turbo::tail_short_mask / tail_long_mask)
    u32 r = __funnelshift_r(w, 0u, h);
    if (has_long) r |= __funnelshift_r(w, 0u, extra);
    return r & 1;
}

// Warp-synchronous confirm of `count` (<= 32) tier-2 survivors c.q2[0..count) (ascending), one lane each.
// ORDERED: every lane collects its candidate's records, a warp prefix sum keeps them in position order and
// they are appended to the warp's own arena segment (no staging, no atomics).
template <bool ORDERED>
__device__ void turbo_drain(const Ctx &c, const uint2 *q2, u64 span_base, u32 count, u32 &n_probe, Rec *seg, u32 &cursor) {
    const u32 idx = lane_id();
    if (ORDERED) {
        Collector col;
        col.n = 0;
        const Ctx cc{c.db, c.cp, c.ar, c.q, c.stage, c.stage_n, c.stage_cap, c.conf, &col};
        if (idx < count) { const uint2 e = q2[idx]; turbo_confirm(cc, span_base + e.x, e.y, n_probe); }
        const u32 n = col.n < 4 ? col.n : 4;
        if (__any_sync(0xffffffffu, n != 0)) {
            u32 incl = n;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const u32 t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane_id() >= (u32)d) incl += t;
            }
            const u32 at = cursor + incl - n;
            for (u32 k = 0; k < n; k++)
                if (at + k < c.ar.seg_cap) seg[at + k] = col.r[k];
            cursor += __shfl_sync(0xffffffffu, incl, 31);
        }
    } else {
        if (idx < count) { const uint2 e = q2[idx]; turbo_confirm(c, span_base + e.x, e.y, n_probe); }
        __syncwarp();
        if (*(volatile u32 *)c.stage_n >= c.stage_cap / 2) flush_stage(c);
    }
}

// Second tier, lane-local: every lane walks the set bits of its own tier-1 mask -- two per round, from both
// ends, so two independent lookups are in flight -- reads the candidate's tail from the warp's staged text
// (shared memory) and keeps the bit when the tail table hits.
// m: bit 31 - j = position j of the lane's row-0 chunk, bit 15 - j = position j of its row-1 chunk.
template <int K>
__device__ __forceinline__ bool turbo_tier2_one(const u32 *tail_tab, u32 word_shift, bool has_long, const u8 *wtext, u32 ln, u32 lz) {
    const u32 pos = ln * TCHUNK + (lz & 15) + (lz >= 16 ? TROW_BYTES : 0);
    if (K <= 3) {
        // the four bytes ending at the position: K-tail on top, the byte before it right below
        const u32 first = THALO + pos - 3;
        const u32 *wp = reinterpret_cast<const u32 *>(wtext + (first & ~3u));
        const u32 w4 = __funnelshift_r(wp[0], wp[1], first << 3);
        return turbo_tail_hit(tail_tab, word_shift, has_long, w4 >> (8 * (4 - K)), w4 >> (8 * (3 - K)));
    } else {
        const u32 first = THALO + pos - 4;
        const u32 *wp = reinterpret_cast<const u32 *>(wtext + (first & ~3u));
        const u32 a = wp[0], b = wp[1];
        const u32 extra = __funnelshift_r(a, b, first << 3);          // low byte = text[pos - 4]
        const u32 ktail = __funnelshift_r(a, b, (first << 3) + 8);    // wrong when first & 3 == 3 (shift 32): fixed below
        return turbo_tail_hit(tail_tab, word_shift, has_long, (first & 3) == 3 ? b : ktail, extra);
    }
}
template <int K>
__device__ __forceinline__ u32 turbo_tier2(const u32 *tail_tab, const DevDb &db, const u8 *wtext, u32 m) {
    const u32 ln = lane_id(), word_shift = 37 - db.tBL;
    const bool has_long = db.tL != 0;
    u32 keep = 0;
    while (m) {
        const u32 lzA = __clz(m), bitA = 0x80000000u >> lzA;
        m ^= bitA;
        const u32 bitB = m & (0u - m);  // lowest remaining bit (0 when none)
        m ^= bitB;
        const u32 lzB = __clz(bitB | 1u);  // bitB == 0 probes the lane's last position; the result is dropped
        const bool hitA = turbo_tier2_one<K>(tail_tab, word_shift, has_long, wtext, ln, lzA);
        const bool hitB = turbo_tier2_one<K>(tail_tab, word_shift, has_long, wtext, ln, lzB);
        if (hitA) keep |= bitA;
        if (hitB) keep |= bitB;
    }
    return keep;
}

// Warp-synchronous: append the step's tier-2 survivors (mask layout as above) to q2 in position order -- row 0
// lane-major, then row 1 -- as span-relative positions; q2 is confirmed whenever it holds 32.
template <int K, bool ORDERED>
__device__ __forceinline__ void turbo_push(const Ctx &c, u32 m, const u8 *wtext, uint2 *q2, u32 &q2n, u64 span_base, u32 step_pos,
                                           u32 &n_probe, Rec *seg, u32 &cursor) {
    const u32 ln = lane_id(), lt = (1u << ln) - 1;
    const u32 n0 = __popc(m >> 16), n1 = __popc(m & 0xffffu);
    u32 slot0, slot1, total;
    if (!__any_sync(0xffffffffu, (n0 | n1) > 1)) {
        // common case: at most one survivor per lane and row -- two ballots rank them
        const u32 b0 = __ballot_sync(0xffffffffu, n0 != 0), b1 = __ballot_sync(0xffffffffu, n1 != 0);
        slot0 = __popc(b0 & lt);
        slot1 = __popc(b0) + __popc(b1 & lt);
        total = __popc(b0) + __popc(b1);
    } else {
        // one inclusive scan serves both rows (16-bit fields)
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
    // q2 holds fewer than 32 entries on entry; a pass adds at most what fits below TQ2CAP
    u32 done = 0;  // survivors of this step already moved to q2 (warp-uniform)
    while (done < total) {
        const u32 room = TQ2CAP - q2n, lim = done + room;  // this pass takes ranks [done, lim)
        u32 mm = m, s0 = slot0, s1 = slot1;
        while (mm) {
            const u32 lz = __clz(mm);
            mm &= ~(0x80000000u >> lz);
            const bool second = lz >= 16;
            const u32 rank = second ? s1++ : s0++;
            if (rank >= done && rank < lim) {
                const u32 pos = ln * TCHUNK + (lz & 15) + (second ? TROW_BYTES : 0);
                const u32 first = THALO + pos - 3;  // the four bytes ending at the position, K-tail on top
                const u32 *wp = reinterpret_cast<const u32 *>(wtext + (first & ~3u));
                const u32 w4 = __funnelshift_r(wp[0], wp[1], first << 3);
                q2[q2n + rank - done] = make_uint2(step_pos + pos, w4 >> (8 * (4 - K)));
            }
        }
        const u32 took = total - done < room ? total - done : room;
        q2n += took;
        done += took;
        __syncwarp();
        while (q2n >= 32) {
            turbo_drain<ORDERED>(c, q2, span_base, 32, n_probe, seg, cursor);
            const u32 rem = q2n - 32;
            // slide the remainder (at most TQ2CAP - 32 = 32 entries) down
            const uint2 k0 = ln < rem ? q2[32 + ln] : make_uint2(0, 0);
            __syncwarp();
            if (ln < rem) q2[ln] = k0;
            __syncwarp();
            q2n = rem;
        }
    }
}

// register whose low byte is extended-chunk byte `e` (E = {8 bytes before the chunk, the chunk})
template <int e>
__device__ __forceinline__ u32 byte_reg(const u32 (&E)[TEW]) {
    if ((e & 3) == 0) return E[e >> 2];
    return __funnelshift_r(E[e >> 2], (e >> 2) + 1 < TEW ? E[(e >> 2) + 1] : 0u, 8 * (e & 3));
}

// K-1 shared bytes of the pair ending at chunk byte i, packed as turbo::pack_shared packs them.
// The packing puts text[i-2] in the low byte, so x(i) doubles as the shift-amount register for the B test
// of pair i-2 and (K == 4) the A test of pair i+2: a funnel shift only looks at the low 5 bits.
template <int K, int i>
__device__ __forceinline__ u32 shared_reg(const u32 (&E)[TEW]) {
    if (K < 2) return 0;
    constexpr int eS = i - K + 1 + 8, wS = eS >> 2, o = eS & 3;
    constexpr u32 sel = K == 4 ? (u32)((o + 1) | (o << 4) | ((o + 2) << 8) | ((o + 2) << 12))
                      : K == 3 ? (u32)(o | ((o + 1) << 4) | (o << 8) | ((o + 1) << 12))
                               : (u32)(o * 0x1111);
    return __byte_perm(E[wS], wS + 1 < TEW ? E[wS + 1] : 0u, sel);
}

// Pairs (i-1, i) for i = 1, 3, ..., TCHUNK-1, unrolled by recursion.  xi = shared_reg<K, i>, xprev =
// shared_reg<K, i-2>; the candidate mask is built first-position-on-top.
template <int K, int i>
struct TurboPairs {
    static __device__ __forceinline__ void run(const u64 *tab, u32 H1shift, const u32 (&E)[TEW], u32 xprev, u32 xi, u32 &acc) {
        u64 ent;
        if (K == 3) {  // direct 7 + 7 bit index (turbo::pair_row), as a byte offset: two shifts and one LOP3
            const u32 off = ((xi << 3) & 0x3f8u) | ((xi << 2) & 0x1fc00u);
            ent = *reinterpret_cast<const u64 *>(reinterpret_cast<const u8 *>(tab) + off);
        } else {
            ent = tab[(xi * turbo::PAIR_MULT) >> H1shift];
        }
        const u32 xnext = i + 2 < TCHUNK ? shared_reg<K, (i + 2 < TCHUNK ? i + 2 : i)>(E) : 0u;
        u32 yA, yB;
        if (K == 4) yA = xprev; else yA = byte_reg<i - K + 8>(E);                         // low 5 bits of text[i-K]
        if ((K == 4 || K == 3) && i + 2 < TCHUNK) yB = xnext; else yB = byte_reg<i + 8>(E);  // text[i]
        // the table stores each 32-bit set bit-reversed: shifting left by low5(byte) brings the flag to bit 31,
        // and a funnel shift by one appends it to the candidate mask (first position ends up in the top bit)
        const u32 ra = __funnelshift_l(0u, (u32)ent, yA);
        acc = __funnelshift_l(ra, acc, 1);
        const u32 rb = __funnelshift_l(0u, (u32)(ent >> 32), yB);
        acc = __funnelshift_l(rb, acc, 1);
        TurboPairs<K, i + 2>::run(tab, H1shift, E, xi, xnext, acc);
    }
};
template <int K>
struct TurboPairs<K, TCHUNK + 1> {
    static __device__ __forceinline__ void run(const u64 *, u32, const u32 (&)[TEW], u32, u32, u32 &) {}
};

// TCHUNK end positions of one chunk -> candidate mask, bit (TCHUNK - 1 - j) = end position j is a candidate
template <int K>
__device__ __forceinline__ u32 turbo_chunk(const u64 *tab, u32 H1shift, const u32 (&E)[TEW]) {
    u32 acc = 0;
    TurboPairs<K, 1>::run(tab, H1shift, E, shared_reg<K, -1>(E), shared_reg<K, 1>(E), acc);
    return acc;
}

template <int K, bool ORDERED>
__global__ void __launch_bounds__(turbo_warps<ORDERED>() * 32, 1) k_turbo(const DevDb db, const DevCorpus cp, const DevArena ar) {
    constexpr int TW = turbo_warps<ORDERED>(), NT = TW * 32;
    using WS = TurboWarpShared<ORDERED>;
    extern __shared__ __align__(16) u8 smem[];
    u64 *tab = reinterpret_cast<u64 *>(smem);
    const u32 rows = 1u << db.tH1, bloom_words = 1u << (db.tBL - 5);
    u32 *bloom = reinterpret_cast<u32 *>(smem + (size_t)rows * 8);
    WS *wsh = reinterpret_cast<WS *>(smem + (size_t)rows * 8 + (size_t)bloom_words * 4);
    const u32 warp = threadIdx.x >> 5, ln = lane_id();
    WS &ws = wsh[warp];
    // every warp of the grid takes one contiguous span of 1 KiB steps, balanced to within one step
    const u64 steps_total = (cp.total + TSTEP_BYTES - 1) / TSTEP_BYTES;
    const u64 W = (u64)gridDim.x * TW, w = (u64)blockIdx.x * TW + warp;
    const u64 r0 = steps_total * w / W, r1 = steps_total * (w + 1) / W;
    const u64 span_base = r0 * TSTEP_BYTES;
    const u32 nsteps = (u32)(r1 - r0);
    const u8 *gsrc = cp.text + span_base - THALO;  // FRONT_PAD keeps the first halo in bounds, BACK_PAD the last step
    // the warp's first steps are requested before the tables are copied in, so HBM latency overlaps the copy
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
        // every CTA copies the same tables: each starts at a different offset so that the SMs do not all ask
        // the same L2 slice for the same line at the same moment
        const uint4 *src = reinterpret_cast<const uint4 *>(db.pair_tab);
        uint4 *dst = reinterpret_cast<uint4 *>(tab);
        const u32 n1 = rows / 2, rot1 = (u32)(((u64)blockIdx.x * n1) / gridDim.x) & ~31u;
#pragma unroll 4
        for (u32 i = threadIdx.x; i < n1; i += NT) {
            const u32 j = (i + rot1) & (n1 - 1);
            dst[j] = __ldg(src + j);
        }
        const uint4 *bsrc = reinterpret_cast<const uint4 *>(db.bloom);
        uint4 *bdst = reinterpret_cast<uint4 *>(bloom);
        const u32 n2 = bloom_words / 4, rot2 = (u32)(((u64)blockIdx.x * n2) / gridDim.x) & ~31u;
#pragma unroll 4
        for (u32 i = threadIdx.x; i < n2; i += NT) {
            const u32 j = (i + rot2) & (n2 - 1);
            bdst[j] = __ldg(bsrc + j);
        }
    }
    __syncthreads();

    const Ctx c{db, cp, ar, nullptr, ws.stage, &ws.stage_n, ORDERED ? 0u : (u32)TSTAGE, nullptr, nullptr};
    uint2 *q2 = ws.q2;
    const u32 H1shift = 32 - db.tH1;
    u32 n_cand = 0, n_probe = 0;
    Rec *seg = ORDERED ? ar.recs + (size_t)w * ar.seg_cap : nullptr;
    u32 cursor = 0, q2n = 0;
    // lane chunk of row r: buffer bytes [THALO + r * 512 + ln * 16, +16); the 8 bytes before it come from the
    // neighbouring chunk (or the halo) in the same buffer.  The registers of step s + 1 are loaded while step s
    // is still in its second tier, so the next first tier starts without a shared-memory round trip.
    uint4 cur0 = make_uint4(0, 0, 0, 0), cur1 = cur0;
    uint2 prev0 = make_uint2(0, 0), prev1 = prev0;
    if (nsteps) {
        mbar_wait(&ws.mbar[0], 0);
        const u8 *cb = ws.text[0] + THALO + ln * TCHUNK;
        cur0 = *reinterpret_cast<const uint4 *>(cb);
        prev0 = *reinterpret_cast<const uint2 *>(cb - 8);
        cur1 = *reinterpret_cast<const uint4 *>(cb + TROW_BYTES);
        prev1 = *reinterpret_cast<const uint2 *>(cb + TROW_BYTES - 8);
    }
#pragma unroll 1
    for (u32 step = 0; step < nsteps; step++) {
        const u32 b = step % TNBUF;
        const u8 *buf = ws.text[b];
        u32 m;
        {
            const u32 E0[TEW] = {prev0.x, prev0.y, cur0.x, cur0.y, cur0.z, cur0.w};
            m = turbo_chunk<K>(tab, H1shift, E0);
            const u32 E1[TEW] = {prev1.x, prev1.y, cur1.x, cur1.y, cur1.z, cur1.w};
            m = (m << 16) | turbo_chunk<K>(tab, H1shift, E1);
        }
        if (step + 1 < nsteps) {
            const u32 nb = (step + 1) % TNBUF;
            mbar_wait(&ws.mbar[nb], ((step + 1) / TNBUF) & 1);
            const u8 *cb = ws.text[nb] + THALO + ln * TCHUNK;
            cur0 = *reinterpret_cast<const uint4 *>(cb);
            prev0 = *reinterpret_cast<const uint2 *>(cb - 8);
            cur1 = *reinterpret_cast<const uint4 *>(cb + TROW_BYTES);
            prev1 = *reinterpret_cast<const uint2 *>(cb + TROW_BYTES - 8);
        }
        // m: bit 31 - j = position j of the lane's row-0 chunk, bit 15 - j = position j of its row-1 chunk
        const u32 step_pos = step * TSTEP_BYTES;
        {
            const u64 g0 = span_base + step_pos + ln * TCHUNK;  // first byte of the row-0 chunk
            if (g0 + TROW_BYTES + TCHUNK > cp.total) {  // clip end positions beyond the corpus
                const long long left0 = (long long)cp.total - (long long)g0, left1 = left0 - TROW_BYTES;
                const u32 k0 = left0 >= 16 ? 0xffffu : left0 <= 0 ? 0u : (0xffffu << (16 - (int)left0)) & 0xffffu;
                const u32 k1 = left1 >= 16 ? 0xffffu : left1 <= 0 ? 0u : (0xffffu << (16 - (int)left1)) & 0xffffu;
                m &= (k0 << 16) | k1;
            }
        }
        n_cand += __popc(m);
        m = turbo_tier2<K>(bloom, db, buf, m);
        if (__any_sync(0xffffffffu, m != 0)) turbo_push<K, ORDERED>(c, m, buf, q2, q2n, span_base, step_pos, n_probe, seg, cursor);
        __syncwarp();
        // the buffer is free: fetch the step TNBUF ahead into it
        if (ln == 0 && step + TNBUF < nsteps)
            bulk_load(ws.text[b], gsrc + (size_t)(step + TNBUF) * TSTEP_BYTES, TBUF_BYTES, &ws.mbar[b]);
    }
    if (q2n) turbo_drain<ORDERED>(c, q2, span_base, q2n, n_probe, seg, cursor);
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

// Ordered mode epilogue: exclusive scan of the per-warp segment counts, then gather the segments into one
// contiguous, position-ordered array (as Rec and as hs_gpu_match_t), one warp per segment.  ctl[4] = records
// kept, ctl[5] = largest count any warp wanted (> seg_cap means the launch must be repeated with larger
// segments).  Block 0 also clears `ctl_next`, the control block of the next scan.
// A train of scans (several corpora, or the chunks of one host batch, scanned back to back) shares one output array:
// `prev` is the control block of the member before this one (its word 6 = records written so far; nullptr for the
// first), `buf_base` the number of buffers before this member; ctl[6] = records written including this member.  Writes
// beyond `out_cap` are dropped and show as ctl[6] > out_cap (the host grows the array and repeats the train).
__global__ void __launch_bounds__(256) k_compact(const Rec *recs, const u32 *seg_count, u32 seg_cap, u32 W, Rec *out_recs,
                                                 hs_gpu_match_t *out, unsigned long long *ctl, unsigned long long *ctl_next,
                                                 const unsigned long long *prev, u32 buf_base, u64 out_cap) {
    __shared__ u32 pre[8193];
    __shared__ u32 part[256];
    __shared__ u32 wsum[8], wmax[8];
    const u32 per = (W + 255) / 256, t = threadIdx.x, ln = t & 31, wid = t >> 5;
    u32 sum = 0, big = 0;
    for (u32 i = t * per; i < (t + 1) * per && i < W; i++) {
        const u32 cnt = seg_count[i];
        big = cnt > big ? cnt : big;
        sum += cnt < seg_cap ? cnt : seg_cap;
    }
    // block-wide exclusive scan of the 256 partial sums (warp scans + scan of the 8 warp totals)
    u32 incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 v = __shfl_up_sync(0xffffffffu, incl, d);
        if (ln >= (u32)d) incl += v;
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        const u32 v = __shfl_down_sync(0xffffffffu, big, d);
        big = v > big ? v : big;
    }
    if (ln == 31) wsum[wid] = incl;
    if (ln == 0) wmax[wid] = big;
    __syncthreads();
    u32 base = 0, total = 0, m = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        if ((u32)k < wid) base += wsum[k];
        total += wsum[k];
        m = wmax[k] > m ? wmax[k] : m;
    }
    part[t] = base + incl - sum;
    const u64 out_base = prev ? prev[6] : 0ull;
    if (blockIdx.x == 0) {
        if (t == 0) { ctl[4] = total; ctl[5] = m; ctl[6] = out_base + total; }
        if (ctl_next && t < 8) ctl_next[t] = 0;
    }
    __syncthreads();
    u32 run = part[t];
    for (u32 i = t * per; i < (t + 1) * per && i < W; i++) {
        pre[i] = run;
        const u32 cnt = seg_count[i];
        run += cnt < seg_cap ? cnt : seg_cap;
    }
    __syncthreads();
    for (u32 w = blockIdx.x * 8 + wid; w < W; w += gridDim.x * 8) {
        const u32 cnt = seg_count[w], n = cnt < seg_cap ? cnt : seg_cap;
        const u64 at = out_base + pre[w];
        for (u32 i = ln; i < n; i += 32) {
            if (at + i >= out_cap) break;
            Rec r = recs[(size_t)w * seg_cap + i];
            r.buf += buf_base;
            out_recs[at + i] = r;
            out[at + i] = hs_gpu_match_t{r.buf, r.id, (uint64_t)r.to};
        }
    }
}
