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
constexpr int LX_STAGE = 64;  // staged match records per warp

struct DevLimex {
    const u64 *tab;       // limex_dev::Tables blob in HBM
    u32 G, n_exc, shift_count, any_squash;
    u64 shift_amounts;
    u32 o_reach, o_shift, o_succ, o_squash, o_meta, o_masks, o_base, total_words;
    u32 accept_abs, accept_eod_abs, nfa_flags;
};

struct LxWarpShared {
    Rec stage[LX_STAGE];
    u32 stage_n;
    u32 pad[3];
};

__host__ __device__ inline size_t limex_smem_bytes(const DevLimex &lx) {
    return (size_t)lx.total_words * 8 + sizeof(LxWarpShared) * LX_WARPS;
}

__device__ __noinline__ void lx_run_program(const Ctx &c, u32 prog, u32 buf, const u8 *btext, u32 blen, u32 to) {
    run_program(c, prog, buf, btext, blen, to);
}

// fire a MO_INVALID_IDX-terminated list of report programs (limexRunReports -> roseReportAdaptor)
__device__ __noinline__ void lx_run_reports(const Ctx &c, u32 list_abs, u32 buf, const u8 *btext, u32 blen, u32 to) {
    for (u32 k = 0; k < 65536; k++) {
        const u32 prog = ld32(c.db.bc, list_abs + 4 * k);
        if (prog == 0xffffffffu) return;
        lx_run_program(c, prog, buf, btext, blen, to);
    }
}

// accepts of state word `a` (already AND-ed with the accept mask) -- moProcessAccepts
template <int G>
__device__ __noinline__ void lx_accepts(const Ctx &c, u64 a, u64 amask, const u32 *sh_base, u32 w, u32 table_abs,
                                           u32 buf, const u8 *btext, u32 blen, u32 to) {
    // every lane walks its own word: accept entries are independent of one another
    const u32 base = sh_base[w];
    while (a) {
        const u32 bit = (u32)__ffsll((long long)a) - 1;
        a &= a - 1;
        const u32 idx = base + (u32)__popcll(amask & ((1ull << bit) - 1));
        const u32 at = table_abs + idx * (u32)nfaaccept::SIZE;
        const u32 rep = ld32(c.db.bc, at + (u32)nfaaccept::reports);
        if (c.db.bc[at + nfaaccept::single_report]) lx_run_program(c, rep, buf, btext, blen, to);
        else lx_run_reports(c, c.db.lx_abs + rep, buf, btext, blen, to);
    }
}

template <int G>
__global__ void __launch_bounds__(LX_THREADS, 3) k_limex(const DevDb db, const DevLimex lx, const DevCorpus cp,
                                                         const DevArena ar, unsigned long long *next_buf) {
    extern __shared__ __align__(16) u8 smem[];
    u64 *tab = reinterpret_cast<u64 *>(smem);
    LxWarpShared *wsh = reinterpret_cast<LxWarpShared *>(smem + (size_t)lx.total_words * 8);
    for (u32 i = threadIdx.x; i < lx.total_words; i += LX_THREADS) tab[i] = lx.tab[i];
    const u32 warp = threadIdx.x >> 5, ln = lane_id();
    if (ln == 0) wsh[warp].stage_n = 0;
    __syncthreads();

    const u64 *sh_reach = tab + lx.o_reach, *sh_succ = tab + lx.o_succ, *sh_squash = tab + lx.o_squash,
              *sh_meta = tab + lx.o_meta, *sh_masks = tab + lx.o_masks;
    const u32 *sh_base = reinterpret_cast<const u32 *>(tab + lx.o_base);
    const Ctx c{db, cp, ar, nullptr, wsh[warp].stage, &wsh[warp].stage_n, LX_STAGE, nullptr, nullptr};

    const u32 w = ln & (G - 1), gbase = ln & ~(u32)(G - 1);
    const u32 gmask = (G == 32 ? 0xffffffffu : ((1u << G) - 1u)) << gbase;
    const u64 initm = sh_masks[0 * G + w], excm = sh_masks[1 * G + w], accm = sh_masks[2 * G + w],
              eodm = sh_masks[3 * G + w];
    u64 shm[8];
#pragma unroll
    for (int k = 0; k < 8; k++) shm[k] = tab[lx.o_shift + k * G + w];
    const u64 amts = lx.shift_amounts;
    const u32 nshift = lx.shift_count;
    const bool any_squash = lx.any_squash != 0;

    // per-buffer state
    u64 s = 0, c_est = 0, c_succ = 0;
    u32 c_rep = 0xffffffffu;
    bool c_valid = false;
    u32 buf = 0, len = 0, i = 0;
    const u8 *btext = nullptr;
    const uint4 *chunk = nullptr;
    u64 lo = 0, hi = 0;
    u32 in_chunk = 0;  // bytes left in (lo, hi)
    bool have = false;
    unsigned long long n_steps = 0;

    for (;;) {
        if (!have || i == len) {
            if (have) {
                // ---- end of buffer: accepts at offset len, then EOD accepts (runtime.c:273-278) ----
                bool live = s != 0;
                if (G > 1) live = __ballot_sync(gmask, live) != 0;
                if (live) {
                    lx_accepts<G>(c, s & accm, accm, sh_base + 1 * G, w, lx.accept_abs, buf, btext, len, len);
                    if (lx.nfa_flags & nfa::ACCEPTS_EOD)
                        lx_accepts<G>(c, s & eodm, eodm, sh_base + 2 * G, w, lx.accept_eod_abs, buf, btext, len, len);
                }
                have = false;
            }
            // ---- next buffer of this group ----
            unsigned long long nb = 0;
            if (w == 0) nb = atomicAdd(next_buf, 1ull);
            if (G > 1) nb = __shfl_sync(gmask, nb, gbase);
            if (nb >= cp.count) break;
            buf = (u32)nb;
            const u64 o0 = cp.off[buf], o1 = cp.off[buf + 1];
            len = (u32)(o1 - o0);
            if (!len || !passes_gates(db, len)) continue;
            btext = cp.text + o0;
            const uintptr_t ap = reinterpret_cast<uintptr_t>(btext);
            chunk = reinterpret_cast<const uint4 *>(ap & ~(uintptr_t)15);
            const u32 skip = (u32)(ap & 15);
            const uint4 v = *chunk++;
            lo = (u64)v.x | ((u64)v.y << 32);
            hi = (u64)v.z | ((u64)v.w << 32);
            if (skip >= 8) { lo = hi >> (8 * (skip - 8)); hi = 0; }
            else if (skip) { lo = (lo >> (8 * skip)) | (hi << (64 - 8 * skip)); hi >>= 8 * skip; }
            in_chunk = 16 - skip;
            i = 0;
            s = initm;
            c_valid = false;
            have = true;
        }
        // ---- one byte ----
        if (!in_chunk) {
            const uint4 v = *chunk++;
            lo = (u64)v.x | ((u64)v.y << 32);
            hi = (u64)v.z | ((u64)v.w << 32);
            in_chunk = 16;
        }
        const u32 ch = (u32)lo & 0xffu;
        lo = (lo >> 8) | (hi << 56);
        hi >>= 8;
        in_chunk--;
        n_steps++;

        u64 succ = 0;
#pragma unroll
        for (int k = 0; k < 8; k++)
            if (k < (int)nshift) succ |= (s & shm[k]) << ((u32)(amts >> (8 * k)) & 63u);

        const u64 e = s & excm;
        u32 eb;
        if (G > 1) eb = __ballot_sync(gmask, e != 0); else eb = e != 0 ? (1u << ln) : 0u;
        if (eb) {
            bool hit = c_valid && e == c_est;
            if (G > 1) hit = __ballot_sync(gmask, !hit) == 0;
            if (hit) {
                succ |= c_succ;
                if (c_rep != 0xffffffffu && w == 0) lx_run_reports(c, c_rep, buf, btext, len, i);
            } else {
                u64 ls = 0;
                u32 rep_seen = 0xffffffffu;
                bool cacheable = true;
                while (eb) {
                    const u32 src = (u32)__ffs((int)eb) - 1;
                    eb &= eb - 1;
                    const u32 ws = src - gbase;
                    u64 ew = G > 1 ? __shfl_sync(gmask, e, src) : e;
                    const u64 xm = sh_masks[1 * G + ws];
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
                            if (w == 0) lx_run_reports(c, rep, buf, btext, len, i);
                            if (rep_seen == 0xffffffffu || rep_seen == rep) rep_seen = rep; else cacheable = false;
                        }
                    }
                }
                succ |= ls;
                if (cacheable) { c_est = e; c_succ = ls; c_rep = rep_seen; c_valid = true; }
            }
        } else {
            // no exception state is on: if no state at all is on, the automaton is dead for good
            bool dead = s == 0;
            if (G > 1) dead = __ballot_sync(gmask, !dead) == 0;
            if (dead) { i = len; s = 0; continue; }
        }
        s = succ & sh_reach[ch * G + w];
        i++;
    }
    __syncwarp();
    flush_stage(c);
#pragma unroll
    for (int d = 16; d; d >>= 1) n_steps += __shfl_down_sync(0xffffffffu, n_steps, d);
    if (ln == 0 && n_steps) atomicAdd(&ar.stats[0], n_steps / G);
}
