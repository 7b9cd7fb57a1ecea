// limex.cuh -- LimEx NFA stepping on the device (included by hs_gpu.cu; tables: limex_tables.h).
//
// What it computes is the reference's block-mode run of a sole-outfix LimEx (src/runtime.c:246-279 ->
// nfaExecLimEx*_Q, src/nfa/limex_runtime_impl.h:620-714, inner loop :209-244):
//     s = init                                                     (MQE_TOP at 0, TOP_FN)
//     for i in [0, len):  succ  = OR_k ((s & shift[k]) << shiftAmount[k])           (NFA_EXEC_GET_LIM_SUCC)
//                         for every exception state in s: fire its reports at offset i, succ &= squash,
//                                                         succ |= successors         (limex_exceptional.h:92-204)
//                         s = succ & reach[text[i]]
//     accepts of s at offset len (accept table), then EOD accepts (limex_common_impl.h:116-213)
// Acceleration (run_accel) and the exception cache are CPU shortcuts with no effect on the result.
//
// Geometry: a buffer is stepped by a GROUP of G lanes, lane w holding 64-bit word w of the state (G = 1 for
// LimEx32/64, 2 for 128, 4 for 256, 8 for 384/512).  Shifts never cross a 64-bit word (limex_tables.h), so the
// shift half of a step is lane-local; an exception's successor row is read by all G lanes at once (one word
// each).  Groups are persistent and take buffers from a global counter, so lanes of a warp stay busy on
// different buffers whatever the buffer lengths are; all tables sit in shared memory.
#pragma once

constexpr int LX_WARPS = 8;
constexpr int LX_THREADS = LX_WARPS * 32;
constexpr int LX_STAGE = 64;  // staged match records per warp (report kernel)

struct DevLimex {
    const u64 *tab;       // limex_dev::Tables blob in HBM
    u32 G, n_exc, shift_count, any_squash, can_die;
    u32 shmul[8];         // 1 << shiftAmount[k] (amounts <= MAX_SHIFT_AMOUNT = 16), constant-bank operands
    u32 o_reach, o_shift, o_shamt, o_csucc, o_succ, o_squash, o_meta, o_masks, o_base, total_words;
    u32 accept_abs, accept_eod_abs, nfa_flags;
};

// What the stepping kernel hands to the report kernel: "the NFA raised report list / report program `what` at
// offset `to` of buffer `buf`" -- the arguments of the NfaCallback (roseReportAdaptor, src/rose/match.c:611-633).
struct __align__(16) LxEvent {
    u32 buf, to, what, single;  // single: 1 = `what` is one program offset; 0 = absolute offset of a MO_INVALID_IDX-
                                // terminated list (LimEx); 2 = absolute offset of a counted report_list (McClellan)
};
struct DevLxEvents {
    LxEvent *ev;
    unsigned long long *count;  // events wanted (may exceed cap)
    u64 cap;
};

struct LxWarpShared {
    Rec stage[LX_STAGE];
    u32 stage_n;
    u32 pad[3];
};

__host__ __device__ inline size_t limex_smem_bytes(const DevLimex &lx) { return (size_t)lx.total_words * 8; }

__device__ __forceinline__ void lx_event(const DevLxEvents &q, u32 buf, u32 to, u32 what, u32 single) {
    const unsigned long long slot = atomicAdd(q.count, 1ull);
    if (slot < q.cap) q.ev[slot] = LxEvent{buf, to, what, single};
}

// accepts of state word `a` (already AND-ed with the accept mask) -- moProcessAccepts; every lane walks its own
// word: accept entries are independent of one another
__device__ __forceinline__ void lx_accepts(const DevLxEvents &q, const u8 *bc, u32 lx_abs, u64 a, u64 amask, u32 base,
                                           u32 table_abs, u32 buf, u32 to) {
    while (a) {
        const u32 bit = (u32)__ffsll((long long)a) - 1;
        a &= a - 1;
        const u32 idx = base + (u32)__popcll(amask & ((1ull << bit) - 1));
        const u32 at = table_abs + idx * (u32)nfaaccept::SIZE;
        const u32 rep = ld32(bc, at + (u32)nfaaccept::reports);
        const u32 single = bc[at + nfaaccept::single_report];
        lx_event(q, buf, to, single ? rep : lx_abs + rep, single);
    }
}

// One shift of the limited model, amount uniform over the launch: (s & shift[k]) << shiftAmount[k]
// written as a multiplication by 2^amount (amounts are < 32): the two IMADs go to the fma pipe, which this loop
// leaves idle, instead of two more SHFs on the alu pipe that bounds it
#define LX_SHIFT(k) succ |= (s & my_shift[(k) * G]) * (u64)lx.shmul[k]

template <int G>
__global__ void __launch_bounds__(LX_THREADS, 4) k_limex(const DevDb db_, const DevLimex lx, const DevCorpus cp_,
                                                         const DevLxEvents evq, unsigned long long *next_buf) {
    extern __shared__ __align__(16) u8 smem[];
    u64 *tab = reinterpret_cast<u64 *>(smem);
    for (u32 i = threadIdx.x; i < lx.total_words; i += LX_THREADS) tab[i] = lx.tab[i];
    const u32 ln = lane_id();
    __syncthreads();

    const u64 *sh_succ = tab + lx.o_succ, *sh_squash = tab + lx.o_squash, *sh_meta = tab + lx.o_meta,
              *sh_masks = tab + lx.o_masks;
    const u32 *sh_base = reinterpret_cast<const u32 *>(tab + lx.o_base);
    constexpr u32 FULL = 0xffffffffu;
    const u32 w = ln & (G - 1), gbase = ln & ~(u32)(G - 1);
    const u32 gmask = (G == 32 ? FULL : ((1u << G) - 1u)) << gbase;
    const u64 excm = sh_masks[1 * G + w];        // exception states the device still has to look at
    u64 csucc = tab[lx.o_csucc + w];             // successors switched on at every step (limex_tables.h)
    const u32 nshift = lx.shift_count;
    const bool any_squash = lx.any_squash != 0, can_die = lx.can_die != 0;
    const u64 *my_shift = tab + lx.o_shift + w, *my_reach = tab + lx.o_reach + w;
    const u32 count = cp_.count;

    // Per-buffer state.  The exception cache starts as "no exception state on -> nothing".
    // For G > 1 every vote and shuffle below is a FULL-WARP operation at a point all 32 lanes pass once per trip
    // of the loop (groups that ran out of buffers idle along with an empty state until the whole warp is done),
    // so the hardware takes the converged path; each group reads its own bits of the result.
    u64 s = 0, c_est = 0, c_succ = 0;
    u32 c_rep = 0xffffffffu;
    u32 buf = 0, len = 0, i = 0;
    const uint4 *chunk = reinterpret_cast<const uint4 *>(cp_.text);
    u64 lo = 0, hi = 0;
    uint4 nxt = make_uint4(0, 0, 0, 0);
    u32 in_chunk = 0;  // bytes left in (lo, hi)
    bool have = false, done = false;

    for (;;) {
        if (!done && i == len) {
            if (have) {
                // ---- end of buffer: accepts at offset len, then EOD accepts (runtime.c:273-278) ----
                if (s) {
                    const u64 accm = sh_masks[2 * G + w];
                    lx_accepts(evq, db_.bc, db_.lx_abs, s & accm, accm, sh_base[1 * G + w], lx.accept_abs, buf, len);
                    if (lx.nfa_flags & nfa::ACCEPTS_EOD) {
                        const u64 eodm = sh_masks[3 * G + w];
                        lx_accepts(evq, db_.bc, db_.lx_abs, s & eodm, eodm, sh_base[2 * G + w], lx.accept_eod_abs, buf, len);
                    }
                }
                have = false;
            }
            // ---- next buffer of this group (empty and gated-out buffers are skipped here) ----
            s = 0; len = 0; i = 0;
            c_est = 0; c_succ = 0; c_rep = 0xffffffffu;  // an idle group must never look like a cache miss
            for (;;) {
                unsigned long long nb = 0;
                if (w == 0) nb = atomicAdd(next_buf, 1ull);
                if (G > 1) nb = __shfl_sync(gmask, nb, gbase);
                if (nb >= count) { done = true; csucc = 0; break; }  // idle from here on: nothing may switch on
                buf = cp_.order ? cp_.order[nb] : (u32)nb;   // longest buffers first: the tail of the launch is short buffers
                const u64 o0 = cp_.off[buf], o1 = cp_.off[buf + 1];
                const u32 l = (u32)(o1 - o0);
                if (!l || !passes_gates(db_, l)) continue;
                len = l;
                const uintptr_t ap = reinterpret_cast<uintptr_t>(cp_.text + o0);
                chunk = reinterpret_cast<const uint4 *>(ap & ~(uintptr_t)15);
                const u32 skip = (u32)(ap & 15);
                const uint4 v = chunk[0];
                nxt = chunk[1];  // always inside the padded text (BACK_PAD)
                chunk += 2;
                lo = (u64)v.x | ((u64)v.y << 32);
                hi = (u64)v.z | ((u64)v.w << 32);
                if (skip >= 8) { lo = hi >> (8 * (skip - 8)); hi = 0; }
                else if (skip) { lo = (lo >> (8 * skip)) | (hi << (64 - 8 * skip)); hi >>= 8 * skip; }
                in_chunk = 16 - skip;
                s = sh_masks[0 * G + w];
                have = true;
                break;
            }
        }
        if (G > 1) { if (__all_sync(FULL, done)) break; } else if (done) break;

        // ---- one byte ----
        if (!in_chunk) {
            lo = (u64)nxt.x | ((u64)nxt.y << 32);
            hi = (u64)nxt.z | ((u64)nxt.w << 32);
            in_chunk = 16;
            if (i + 16 < len) nxt = *chunk;  // the chunk after this one, if the buffer reaches it
            chunk++;
        }
        const u32 ch = (u32)lo & 0xffu;
        lo = (lo >> 8) | (hi << 56);
        hi >>= 8;
        in_chunk--;

        u64 succ = csucc;
        switch (nshift) {
        case 8: LX_SHIFT(7);  // fallthrough
        case 7: LX_SHIFT(6);  // fallthrough
        case 6: LX_SHIFT(5);  // fallthrough
        case 5: LX_SHIFT(4);  // fallthrough
        case 4: LX_SHIFT(3);  // fallthrough
        case 3: LX_SHIFT(2);  // fallthrough
        case 2: LX_SHIFT(1);  // fallthrough
        case 1: LX_SHIFT(0);  // fallthrough
        default: break;
        }

        const u64 e = s & excm;
        u32 missw;  // lanes (of the whole warp) whose exception word differs from the cached one
        if (G > 1) missw = __ballot_sync(FULL, e != c_est); else missw = e != c_est ? gmask : 0u;
        if (missw) {
            const bool mine = (missw & gmask) != 0;  // my group missed
            u32 ebw;
            if (G > 1) ebw = __ballot_sync(FULL, mine && e != 0); else ebw = e != 0 ? gmask : 0u;
            u32 eb = ebw & gmask;
            u64 ls = 0;
            u32 rep_seen = 0xffffffffu;
            bool cacheable = true;
            while (ebw) {  // one exception word per group and round
                const u32 src = eb ? (u32)__ffs((int)eb) - 1 : ln;
                u64 ew = G > 1 ? __shfl_sync(FULL, e, src) : e;
                if (!eb) ew = 0;
                const u32 ws = src - gbase;
                const u64 xm = sh_masks[4 * G + ws];
                const u32 base = sh_base[ws];
                while (ew) {
                    const u32 bit = (u32)__ffsll((long long)ew) - 1;
                    ew &= ew - 1;
                    const u32 idx = base + (u32)__popcll(xm & ((1ull << bit) - 1));
                    ls |= sh_succ[idx * G + w];
                    const u64 meta = sh_meta[idx];
                    if (any_squash && (meta >> 32)) { succ &= sh_squash[idx * G + w]; cacheable = false; }
                    const u32 rep = (u32)meta;
                    if (rep != 0xffffffffu) {
                        if (w == 0) lx_event(evq, buf, i, rep, 0);
                        if (rep_seen == 0xffffffffu || rep_seen == rep) rep_seen = rep; else cacheable = false;
                    }
                }
                eb &= eb - 1;
                if (G > 1) ebw = __ballot_sync(FULL, eb != 0); else ebw = 0;
            }
            if (mine) {
                succ |= ls;
                if (cacheable) { c_est = e; c_succ = ls; c_rep = rep_seen; }
            } else {
                succ |= c_succ;
                if (c_rep != 0xffffffffu && w == 0) lx_event(evq, buf, i, c_rep, 0);
            }
        } else {
            succ |= c_succ;
            if (c_rep != 0xffffffffu && w == 0) lx_event(evq, buf, i, c_rep, 0);
        }
        s = succ & my_reach[ch * G];
        i++;
        if (can_die) {
            // a state vector with nothing on stays that way (no TOP arrives mid-buffer): skip to the end
            bool nz = s != 0;
            if (G > 1) nz = (__ballot_sync(FULL, nz) & gmask) != 0;
            if (!nz && i < len) i = len;
        }
    }
}
#undef LX_SHIFT

// The report half: one thread per event runs the Rose report program(s) the NFA named (limexRunReports ->
// roseReportAdaptor -> roseRunProgram) and stages the resulting match records.
__global__ void __launch_bounds__(LX_THREADS) k_limex_report(const DevDb db, const DevCorpus cp, const DevArena ar,
                                                             const DevLxEvents evq) {
    __shared__ LxWarpShared wsh[LX_WARPS];
    const u32 warp = threadIdx.x >> 5, ln = lane_id();
    if (ln == 0) wsh[warp].stage_n = 0;
    __syncwarp();
    const Ctx c{db, cp, ar, nullptr, wsh[warp].stage, &wsh[warp].stage_n, LX_STAGE, nullptr, nullptr};
    unsigned long long n = *evq.count;
    if (n > evq.cap) n = evq.cap;  // overflow: the host grows the queue and repeats the scan
    const unsigned long long stride = (unsigned long long)gridDim.x * LX_THREADS;
    for (unsigned long long base = (unsigned long long)blockIdx.x * LX_THREADS + warp * 32; base < n; base += stride) {
        const unsigned long long k = base + ln;
        if (k < n) {
            const LxEvent ev = evq.ev[k];
            const u64 o0 = cp.off[ev.buf];
            const u32 blen = (u32)(cp.off[ev.buf + 1] - o0);
            const u8 *btext = cp.text + o0;
            if (ev.single == 1) {
                run_program(c, ev.what, ev.buf, btext, blen, ev.to);
            } else if (ev.single == 2) {  // counted list: struct report_list {u32 count; ReportID report[]} (McClellan)
                const u32 cnt = ld32(db.bc, ev.what);
                for (u32 j = 0; j < cnt; j++) run_program(c, ld32(db.bc, ev.what + 4 + 4 * j), ev.buf, btext, blen, ev.to);
            } else {
                for (u32 j = 0; j < 65536; j++) {
                    const u32 prog = ld32(db.bc, ev.what + 4 * j);
                    if (prog == 0xffffffffu) break;
                    run_program(c, prog, ev.buf, btext, blen, ev.to);
                }
            }
        }
        flush_stage(c);
    }
}
