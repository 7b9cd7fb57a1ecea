// limex_tables.h -- host side: re-lay a LimExNFA{32..512} (src/nfa/limex_internal.h:114-160) for the device.
//
// The device steps the very automaton the database holds -- same reach sets, same shifts, same exceptions, same
// report lists -- but reads it from tables shaped for a group of G lanes that each own ONE 64-bit word of the
// state vector.  That split is exact, not an approximation: the reference compiler never lets a shift edge cross
// a 64-bit lane of the state ("can't jump over a bollard", src/nfa/limex_compile.cpp:245-255) and the runtime
// shifts each 64-bit lane on its own (lshift64_m128/m256/m512 in LSHIFT_STATE), so the shift half of a step is
// lane-local; only exception successors cross lanes, and those are table rows.
//
// Blob layout (u64 words; offsets in DevLimex):
//   reach[256][G]      reach set of each input byte, reachMap already applied (limex_state_impl.h:64-67)
//   shift[8][G]        shift masks, in the database's order (amounts: Tables::shift_amounts, uniform per launch)
//   shamt[G]           unused
//   csucc[G]           successors switched on at EVERY step: the exception successors of states that are on for
//                      the whole scan (in init, self-loop by a 0-shift, in every byte's reach set, never squashed)
//                      and that neither report nor squash; such states are dropped from masks[1] (the device's
//                      exception mask) but keep their slot in the exception table (ranks use masks[4])
//   exc_succ[n][G]     successors of exception i          (NFAException.successors)
//   exc_squash[n][G]   squash mask of exception i, all-ones when it has none (only present if any squash)
//   exc_meta[n]        lo32 = absolute bytecode offset of the report list or 0xffffffff, hi32 = 1 if it squashes
//   masks[5][G]        init, device exception mask, accept, acceptAtEOD, exceptionMask as in the database
//   base[4][G] (u32)   exceptions / accepts / EOD accepts that precede word w (rank base of its first bit);
//                      [3][w] = number of non-zero shift masks of word w (statistics)
#pragma once
#include "hsb_dbview.h"

#include <vector>

namespace hsb {
namespace limex_dev {

struct Tables {
    u32 G = 1;            // lanes per buffer = 64-bit words per state, padded to a power of two
    u32 words = 1;        // words actually holding state bits
    u32 n_exc = 0, shift_count = 0;
    u64 shift_amounts = 0;  // byte k = shiftAmount[k]
    bool any_squash = false;
    bool can_die = true;    // false when some state is on for the whole scan (then the state vector never empties)
    u32 o_reach = 0, o_shift = 0, o_shamt = 0, o_csucc = 0, o_succ = 0, o_squash = 0, o_meta = 0, o_masks = 0, o_base = 0, total_words = 0;
    u32 accept_abs = 0, accept_eod_abs = 0, limex_abs = 0;  // absolute bytecode offsets
    u32 nfa_flags = 0;
    std::vector<u64> blob;
};

inline u64 state_word(const u8 *p, u32 bytes, u32 w) {  // word w of a `bytes`-byte little-endian bit vector
    u64 v = 0;
    for (u32 b = 0; b < 8; b++) {
        const u32 at = 8 * w + b;
        if (at < bytes) v |= (u64)p[at] << (8 * b);
    }
    return v;
}

inline bool build(const DbView &d, Tables &t) {
    if (d.kind != ENG_LIMEX) return false;
    const limex::Layout L = limex::layout(d.limex_type);
    const u8 *lx = d.bc + d.limex_off;
    t.words = (L.bytes + 7) / 8;
    t.G = 1;
    while (t.G < t.words) t.G *= 2;
    const u32 G = t.G;
    t.limex_abs = d.limex_off;
    t.nfa_flags = d.nfa_flags;
    t.n_exc = rd32(lx, limex::exceptionCount);
    t.shift_count = rd32(lx, L.shiftCount);
    t.shift_amounts = 0;
    for (u32 k = 0; k < t.shift_count; k++) t.shift_amounts |= (u64)rd8(lx, L.shiftAmount + k) << (8 * k);
    t.accept_abs = d.limex_off + rd32(lx, limex::acceptOffset);
    t.accept_eod_abs = d.limex_off + rd32(lx, limex::acceptEodOffset);
    const u8 *exc = lx + rd32(lx, limex::exceptionOffset);
    t.any_squash = false;
    for (u32 i = 0; i < t.n_exc; i++)
        if (rd8(exc + (size_t)i * L.E_SIZE, L.e_hasSquash) != limex::SQUASH_NONE) t.any_squash = true;

    u32 at = 0;
    t.o_reach = at; at += 256 * G;
    t.o_shift = at; at += 8 * G;
    t.o_shamt = at; at += G;
    t.o_csucc = at; at += G;
    t.o_succ = at; at += t.n_exc * G;
    t.o_squash = at; at += t.any_squash ? t.n_exc * G : 0;
    t.o_meta = at; at += t.n_exc;
    t.o_masks = at; at += 5 * G;
    t.o_base = at; at += (4 * G + 1) / 2;
    t.total_words = (at + 1) & ~1u;  // the per-warp staging that follows in shared memory is 16-byte aligned
    t.blob.assign(t.total_words, 0);
    u64 *B = t.blob.data();

    const u8 *reach = lx + L.SIZE;
    for (u32 c = 0; c < 256; c++) {
        const u8 *row = reach + (size_t)rd8(lx, limex::reachMap + c) * L.bytes;
        for (u32 w = 0; w < t.words; w++) B[t.o_reach + c * G + w] = state_word(row, L.bytes, w);
    }
    u32 *base = reinterpret_cast<u32 *>(B + t.o_base);
    std::vector<u64> loop0(G, 0);  // states with a self loop through a 0-shift
    for (u32 w = 0; w < t.words; w++) {
        u32 n = 0;
        for (u32 k = 0; k < t.shift_count; k++) {
            const u64 m = state_word(lx + L.shift + (size_t)k * L.bytes, L.bytes, w);
            const u32 amt = rd8(lx, L.shiftAmount + k);
            if (!m) continue;
            if (!amt) loop0[w] |= m;
            B[t.o_shift + k * G + w] = m;
            n++;
        }
        base[3 * G + w] = n;
    }
    for (u32 i = 0; i < t.n_exc; i++) {
        const u8 *e = exc + (size_t)i * L.E_SIZE;
        const bool sq = rd8(e, L.e_hasSquash) != limex::SQUASH_NONE;
        for (u32 w = 0; w < t.words; w++) {
            B[t.o_succ + i * G + w] = state_word(e + L.e_successors, L.bytes, w);
            if (t.any_squash) B[t.o_squash + i * G + w] = sq ? state_word(e + L.e_squash, L.bytes, w) : ~0ull;
        }
        const u32 rep = rd32(e, L.e_reports);
        B[t.o_meta + i] = (u64)(rep == 0xffffffffu ? rep : d.limex_off + rep) | ((u64)(sq ? 1 : 0) << 32);
    }
    const size_t mask_at[5] = {L.init, L.exceptionMask, L.accept, L.acceptAtEOD, L.exceptionMask};
    for (u32 m = 0; m < 5; m++)
        for (u32 w = 0; w < t.words; w++) B[t.o_masks + m * G + w] = state_word(lx + mask_at[m], L.bytes, w);
    for (u32 m = 0; m < 3; m++) {
        u32 run = 0;
        for (u32 w = 0; w < G; w++) {
            base[m * G + w] = run;
            run += (u32)__builtin_popcountll(B[t.o_masks + (m == 0 ? 4 : m + 1) * G + w]);
        }
    }
    // states that are on from the first step to the last: fold their plain exception successors into csucc
    for (u32 w = 0; w < t.words; w++) {
        u64 on = B[t.o_masks + 0 * G + w] & loop0[w];
        for (u32 c = 0; c < 256; c++) on &= B[t.o_reach + c * G + w];
        if (t.any_squash)
            for (u32 i = 0; i < t.n_exc; i++) on &= B[t.o_squash + i * G + w];
        if (on) t.can_die = false;
        const u64 xm = B[t.o_masks + 4 * G + w];
        for (u64 x = on & xm; x; x &= x - 1) {
            const u32 bit = (u32)__builtin_ctzll(x);
            const u32 idx = base[0 * G + w] + (u32)__builtin_popcountll(xm & ((1ull << bit) - 1));
            const u64 meta = B[t.o_meta + idx];
            if ((u32)meta != 0xffffffffu || (meta >> 32)) continue;  // reports or squashes: stays an exception
            for (u32 v = 0; v < t.words; v++) B[t.o_csucc + v] |= B[t.o_succ + idx * G + v];
            B[t.o_masks + 1 * G + w] &= ~(1ull << bit);
        }
    }
    return true;
}

}  // namespace limex_dev
}  // namespace hsb
