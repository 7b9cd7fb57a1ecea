// mcclellan.cuh -- McClellan DFA stepping on the device (included by hs_gpu.cu; census: hsb_dbview.h).
//
// What it computes is the reference's block-mode run of a sole-outfix McClellan engine (src/runtime.c:246-279 ->
// nfaExecMcClellan{8,16}_Q -> mcclellanExec{8,16}_i, src/nfa/mcclellan.c:227-366 / :445-565):
//     s = start_anchored                                           (MQE_TOP at offset 0, mcclellan.c:740-745)
//     for i in [0, len):  c' = remap[text[i]]
//                         16-bit: e = s < sherman_limit ? succ[(s << alphaShift) + c'] : sherman(s, c')
//                                 e & ACCEPT_FLAG -> reports of state e at offset i + 1;  s = e & STATE_MASK
//                          8-bit: s = succ[(s << alphaShift) + c'];  s >= accept_limit_8 -> reports at offset i + 1
//                         s == 0 (dead) ends the scan
//     NFA_ACCEPTS_EOD: aux[s].accept_eod reports at offset len     (runtime.c:273-278, mcclellanCheckEOD)
// Reports: MCCLELLAN_FLAG_SINGLE -> arb_report, else the counted report list at aux[s].accept (doComplexReport,
// mcclellan.c:43-89; its one-entry cache is a CPU shortcut).  Acceleration (run_mcclellan_accel) skips bytes that
// cannot change the state -- result-neutral, not used here; ACCEL_FLAG bits in table entries are masked off.
//
// Geometry: a DFA step is one dependent table read per byte, so the unit of parallelism is the BUFFER: one
// thread per buffer, threads persistent (first buffer by thread index, later ones from a global counter; buffers of
// any length mix in a warp).  The successor table and a bank-replicated symbol table sit in shared memory when
// they fit (<= MCC_SMEM_MAX), else the successor table is read through L1/L2.  Text is read in aligned 16-byte
// chunks, three chunks ahead; bytes of a chunk outside the buffer freeze the state (predicated read), so all 32
// lanes run the same 16 unrolled steps.  A Sheng engine is re-laid at upload (hs_gpu.cu build_sheng_tables) into the
// 8-bit form of this table and stepped by the same kernel.
// Hot loop = 16 branch-free steps that only remember whether an accept was passed; the rare chunk that passed one
// is stepped again from its start state by the exact loop, which raises the events.
#pragma once

constexpr int MCC_THREADS = 1024;
constexpr size_t MCC_SMEM_MAX = 226 * 1024;  // of the 227 KB a CTA may have
constexpr size_t MCC_RM_BYTES = 256 * 32 * 4;  // symbol table, replicated per bank

struct DevMcc {
    u32 tab_abs;        // absolute bytecode offset of the successor table
    u32 tab_bytes;      // its size (rows below sherman_limit only for 16-bit engines)
    u32 remap_abs;      // absolute bytecode offset of remap[256]
    u32 nfa_abs;        // absolute bytecode offset of struct NFA (report-list and aux offsets are relative to it)
    u32 aux_abs;        // absolute bytecode offset of mstate_aux[0]
    u32 sherman_abs;    // absolute bytecode offset of the first Sherman state
    u32 wide16, as, start, accept_limit_8, sherman_limit, single, arb_report, accepts_eod, in_smem, has_sherman;
};

// Several engines of one database in ONE launch (full-Rose databases of outfix engines): their tables sit side by side
// in shared memory and a work item is a (buffer, engine) pair, so a launch has `n` times the items -- with one buffer
// per thread at most (a 64 MiB batch of ~800-byte payloads has fewer buffers than the grid has threads) that is what
// fills the machine.  Only engines the branch-free loop serves (table in shared memory, no Sherman states) are packed.
constexpr u32 MCC_STREAM_CLOSE = 0x80000000u;   // stream mode, entry state: "no data, raise the end-of-data accepts of this state"
constexpr int MCC_PACK_MAX = 4;
struct DevMccPack {
    DevMcc m[MCC_PACK_MAX];
    u32 tab_off[MCC_PACK_MAX];  // shared-memory offset of each engine's successor table (behind the n symbol tables)
    u32 n;
};

__host__ __device__ inline size_t mcc_smem_bytes(const DevMcc &m) { return MCC_RM_BYTES + (m.in_smem ? (size_t)m.tab_bytes + 8 : 0); }

// doSherman16 (mcclellan_common_impl.h:54-84): a Sherman state lists up to 9 (symbol, successor) exceptions to
// its daddy's row
__device__ __noinline__ u32 mcc_sherman(const u8 *bc, const DevMcc &m, const u16 *tab16, u32 s, u32 cp) {
    const u8 *sh = bc + m.sherman_abs + (size_t)mcclellan::SHERMAN_FIXED_SIZE * (s - m.sherman_limit);
    const u32 len = sh[mcclellan::SHERMAN_LEN_OFFSET];
    for (u32 k = 0; k < len; k++)
        if (sh[mcclellan::SHERMAN_CHARS_OFFSET + k] == cp) {
            const u8 *p = sh + mcclellan::SHERMAN_CHARS_OFFSET + len + 2 * k;
            return (u32)p[0] | ((u32)p[1] << 8);
        }
    const u32 daddy = (u32)sh[mcclellan::SHERMAN_DADDY_OFFSET] | ((u32)sh[mcclellan::SHERMAN_DADDY_OFFSET + 1] << 8);
    return tab16[(daddy << m.as) + cp];
}

// "state `st` accepts at offset `to`": hand the report (list) to the report kernel
__device__ __noinline__ void mcc_accept(const DevLxEvents &q, const u8 *bc, const DevMcc &m, u32 st, u32 buf, u32 to, bool eod) {
    if (m.single && !eod) { lx_event(q, buf, to, m.arb_report, 1); return; }
    const u32 a = m.aux_abs + st * (u32)mcclellan::aux::SIZE;
    const u32 rl = ld32(bc, a + (u32)(eod ? mcclellan::aux::accept_eod : mcclellan::aux::accept));
    if (rl) lx_event(q, buf, to, m.nfa_abs + rl, 2);
}

// One step of the reference's loop, any engine flavour; returns the raw table entry (16-bit: flags included).
template <bool WIDE16>
__device__ __forceinline__ u32 mcc_step(const u8 *bc, const DevMcc &m, const u8 *tab, u32 rowbytes, u32 s, u32 sym) {
    if (!WIDE16) return tab[s * rowbytes + sym];
    if (s < m.sherman_limit) return *reinterpret_cast<const u16 *>(tab + s * rowbytes + sym);
    return mcc_sherman(bc, m, reinterpret_cast<const u16 *>(tab), s, sym >> 1);
}

// SHERMAN = false: every state has a table row, so a chunk is 16 branch-free steps (extract, remap, multiply-add,
// table read, select) that only OR the entries together; a chunk that saw an accept flag (rare) is then stepped
// again from its saved start state by the exact loop below, which raises the events.  SHERMAN = true (16-bit
// engines with Sherman states): the exact loop is the only loop.
template <bool WIDE16, bool SMEM, bool SHERMAN>
__global__ void __launch_bounds__(MCC_THREADS, 1) k_mcclellan(const DevDb db_, const DevMcc m, const DevCorpus cp_,
                                                              const DevLxEvents evq, unsigned long long *next_buf) {
    extern __shared__ __align__(16) u8 smem[];
    // symbol table: rm[c] = remap[c] * (entry size) + (shared-memory address of the table), so that row offset + rm[c]
    // is the address of the entry; one copy per bank (word c * 32 + lane): a warp's 32 lookups never conflict
    u32 *rm_w = reinterpret_cast<u32 *>(smem);
    const u32 tab_sh = SMEM ? (u32)__cvta_generic_to_shared(smem + MCC_RM_BYTES + (m.tab_abs & 3u)) : 0u;
    for (u32 i = threadIdx.x; i < 256 * 32; i += blockDim.x)
        rm_w[i] = (u32)db_.bc[m.remap_abs + (i >> 5)] * (WIDE16 ? 2u : 1u) + tab_sh;
    if (SMEM) {
        const u32 head = m.tab_abs & 3u;  // copy whole words; the table is only 2-aligned in the bytecode (372 = 64 + 308)
        const u32 words = (m.tab_bytes + head + 3) / 4;
        const u32 *src = reinterpret_cast<const u32 *>(db_.bc + (m.tab_abs - head));
        u32 *dst = reinterpret_cast<u32 *>(smem + MCC_RM_BYTES);
        for (u32 i = threadIdx.x; i < words; i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const u32 *rm = rm_w + (threadIdx.x & 31u);
    const u8 *tab = SMEM ? smem + MCC_RM_BYTES + (m.tab_abs & 3u) : db_.bc + m.tab_abs;
    const u32 rowbytes = (WIDE16 ? 2u : 1u) << m.as, count = cp_.count;
    const u32 accept_limit = m.accept_limit_8;
    const u32 lane4 = (threadIdx.x & 31u) * 4u;

    u32 s = 0, buf = 0, len = 0;
    long long i = 0;           // buffer offset of byte 0 of the current chunk (negative for the first chunk)
    const uint4 *chunk = nullptr;
    uint4 a[4];                // this chunk and the three after it; indexed statically (the trip below is unrolled
                               // four times), so a chunk is loaded into the registers its predecessor just left and
                               // is not touched again until it is stepped, 48 steps later
    a[0] = a[1] = a[2] = a[3] = make_uint4(0, 0, 0, 0);
    bool have = false, first = true;

    for (;;) {
        if (!have) {
            // ---- next buffer of this thread (empty and gated-out buffers report nothing: runtime.c:343-371) ----
            unsigned long long nb;
            u32 l = 0;
            u64 o0 = 0;
            for (;;) {
                if (first) {
                    // first buffer: by thread index, no counter traffic (one same-address atomic per thread would
                    // serialise at the L2: ~1 per clock, i.e. longer than the scan itself for small buffers)
                    // warps interleaved across the CTAs, so a batch smaller than the grid still uses every SM
                    nb = ((unsigned long long)(threadIdx.x >> 5) * gridDim.x + blockIdx.x) * 32u + (threadIdx.x & 31u);
                    first = false;
                } else {
                    // later buffers: one atomic per group of lanes that arrive here together
                    const u32 act = __activemask(), ln = threadIdx.x & 31u, leader = (u32)__ffs((int)act) - 1u;
                    unsigned long long base = 0;
                    if (ln == leader) base = atomicAdd(next_buf, (unsigned long long)__popc(act));
                    base = __shfl_sync(act, base, leader);
                    nb = base + (unsigned long long)gridDim.x * blockDim.x + (u32)__popc(act & ((1u << ln) - 1u));
                }
                if (nb >= count) break;
                const u32 b = cp_.order ? cp_.order[nb] : (u32)nb;
                o0 = cp_.off[b];
                l = (u32)(cp_.off[b + 1] - o0);
                if (cp_.init_state && (cp_.init_state[b] & MCC_STREAM_CLOSE)) {
                    // the stream is being closed: nfaCheckFinalState of the state it stands in, at its end (offset 0 of
                    // this data-less block)
                    const u32 cs = cp_.init_state[b] & ~MCC_STREAM_CLOSE;
                    if (cs && m.accepts_eod) mcc_accept(evq, db_.bc, m, cs, b, 0, true);
                    continue;
                }
                if (l && passes_gates(db_, l)) break;
            }
            if (nb >= count) break;
            buf = cp_.order ? cp_.order[nb] : (u32)nb;
            len = l;
            const uintptr_t ap = reinterpret_cast<uintptr_t>(cp_.text + o0);
            chunk = reinterpret_cast<const uint4 *>(ap & ~(uintptr_t)15);
            i = -(long long)(ap & 15);
            a[0] = chunk[0];             // over-reads stay inside the padded text (BACK_PAD)
            a[1] = chunk[1]; a[2] = chunk[2]; a[3] = chunk[3];
            chunk += 4;
            s = cp_.init_state ? cp_.init_state[buf] : m.start;   // stream mode: the state the stream has come to
            have = s != 0;                                        // (a dead stream stays dead)
        }
#pragma unroll
      for (int q = 0; q < 4; q++) {
        if (!have) break;  // the buffer ended in an earlier quarter of this trip
        // ---- one aligned 16-byte chunk: bytes [j0, j1) of it belong to the buffer ----
        const uint4 v = a[q];
        const long long rem = (long long)len - i;  // bytes from chunk byte 0 to the end of the buffer
        // three chunks (48 steps) of loads in flight while this one is stepped: a step is ~40 cycles of dependent
        // latency, a miss to HBM ~1000.  (A rotating FIFO -- cur = q1; q1 = q2; q2 = q3 -- makes the move wait for
        // the load issued one trip earlier: 20 % of the stall samples sat there.)
        if (rem > 64) a[q] = *chunk;
        chunk++;
        const u32 j0 = i < 0 ? (u32)(-i) : 0u;
        const u32 j1 = rem < 16 ? (u32)rem : 16u;
        const u32 valid = ((1u << j1) - 1u) & ~((1u << j0) - 1u);
        const u32 wv[4] = {v.x, v.y, v.z, v.w};
        const u32 s0 = s;
        bool exact = SHERMAN;
        if (!SHERMAN) {
            u32 seen = 0;  // 16-bit: OR of the entries taken; 8-bit: highest state entered
#pragma unroll
            for (int j = 0; j < 16; j++) {
                // byte j of the chunk, times 128 (one 32-word row of the symbol table per byte value), plus my bank
                const u32 w = wv[j >> 2], k = 8 * (j & 3);
                const u32 ra = ((k >= 7 ? w >> (k - 7) : w << (7 - k)) & 0x7f80u) | lane4;
                const u32 sym = *reinterpret_cast<const u32 *>(smem + ra);
                u32 e;
                if (SMEM) {
                    // row offset + (symbol offset + table address): the second term is off the dependent chain, which
                    // is then mask -> multiply-add -> shared-memory read
                    u32 addr;  // written as PTX so the compiler keeps the association
                    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(addr) : "r"(WIDE16 ? (s & mcclellan::STATE_MASK) : s), "r"(rowbytes), "r"(sym));
                    if (WIDE16) { unsigned short t; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(t) : "r"(addr)); e = t; }
                    else asm volatile("ld.shared.u8 %0, [%1];" : "=r"(e) : "r"(addr));
                } else if (WIDE16) e = *reinterpret_cast<const u16 *>(tab + (s & mcclellan::STATE_MASK) * rowbytes + sym);
                else e = tab[s * rowbytes + sym];
                s = ((valid >> j) & 1u) ? e : s;   // bytes outside the buffer freeze the state (flags of s stay as they were)
                if (WIDE16) seen |= s; else seen = max(seen, s);
            }
            if (WIDE16) { exact = (seen & mcclellan::ACCEPT_FLAG) != 0; s &= mcclellan::STATE_MASK; }
            else exact = seen >= accept_limit;
        }
        if (exact) {
            // the reference's loop, byte by byte, raising the accepts (mcclellan.c:267-289 / :482-506)
            s = s0;
            for (u32 j = j0; j < j1 && s; j++) {
                const u32 w = j < 8 ? (j < 4 ? v.x : v.y) : (j < 12 ? v.z : v.w);
                const u32 e = mcc_step<WIDE16>(db_.bc, m, tab, rowbytes, s, rm[((w >> (8 * (j & 3))) & 0xffu) * 32] - tab_sh);
                s = WIDE16 ? (e & mcclellan::STATE_MASK) : e;
                if (WIDE16 ? (e & mcclellan::ACCEPT_FLAG) != 0 : e >= accept_limit)
                    mcc_accept(evq, db_.bc, m, s, buf, (u32)(i + j + 1), false);
            }
        }
        i += 16;
        if (i >= (long long)len || !s) {
            // ---- end of buffer (or dead: nothing can switch on again, mcclellan.c:240-243) ----
            if (cp_.final_state) cp_.final_state[buf] = s;        // stream mode: the data goes on, no end-of-data check
            else if (s && m.accepts_eod) mcc_accept(evq, db_.bc, m, s, buf, len, true);
            have = false;
        }
      }
    }
}

// k_mcclellan for a pack of engines: same loop; the engine is part of the work item, so the symbol-table base, the row
// size and the accept limit are per-thread registers instead of launch constants.
template <bool WIDE16>
__global__ void __launch_bounds__(MCC_THREADS, 1) k_mcclellan_pack(const DevDb db_, const DevMccPack pk, const DevCorpus cp_,
                                                                   const DevLxEvents evq, unsigned long long *next_buf) {
    constexpr bool SMEM = true, SHERMAN = false;
    extern __shared__ __align__(16) u8 smem[];
    // per engine e: symbol table rm_e[c] = remap_e[c] * (entry size) + (shared-memory address of table e), one copy per
    // bank, at smem + e * MCC_RM_BYTES; the successor tables follow the n symbol tables
    const u32 npk = pk.n;
    for (u32 e = 0; e < npk; e++) {
        const DevMcc &me = pk.m[e];
        u32 *rm_w = reinterpret_cast<u32 *>(smem + (size_t)e * MCC_RM_BYTES);
        const u32 tab_sh_e = (u32)__cvta_generic_to_shared(smem + pk.tab_off[e] + (me.tab_abs & 3u));
        for (u32 i = threadIdx.x; i < 256 * 32; i += blockDim.x)
            rm_w[i] = (u32)db_.bc[me.remap_abs + (i >> 5)] * (WIDE16 ? 2u : 1u) + tab_sh_e;
        const u32 head = me.tab_abs & 3u;
        const u32 words = (me.tab_bytes + head + 3) / 4;
        const u32 *src = reinterpret_cast<const u32 *>(db_.bc + (me.tab_abs - head));
        u32 *dst = reinterpret_cast<u32 *>(smem + pk.tab_off[e]);
        for (u32 i = threadIdx.x; i < words; i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const u32 count = cp_.count;
    const unsigned long long items = (unsigned long long)count * npk;
    const u32 lane4 = (threadIdx.x & 31u) * 4u;
    // the current item's engine
    u32 eng = 0, rowbytes = 0, accept_limit = 0, tab_sh = 0;
    const u8 *rm_base = smem;   // symbol table of the item's engine
    const u8 *tab = smem;

    u32 s = 0, buf = 0, len = 0;
    long long i = 0;           // buffer offset of byte 0 of the current chunk (negative for the first chunk)
    const uint4 *chunk = nullptr;
    uint4 a[4];                // this chunk and the three after it; indexed statically (the trip below is unrolled
                               // four times), so a chunk is loaded into the registers its predecessor just left and
                               // is not touched again until it is stepped, 48 steps later
    a[0] = a[1] = a[2] = a[3] = make_uint4(0, 0, 0, 0);
    bool have = false, first = true;

    for (;;) {
        if (!have) {
            // ---- next buffer of this thread (empty and gated-out buffers report nothing: runtime.c:343-371) ----
            unsigned long long nb;
            u32 l = 0;
            u64 o0 = 0;
            for (;;) {
                if (first) {
                    // first buffer: by thread index, no counter traffic (one same-address atomic per thread would
                    // serialise at the L2: ~1 per clock, i.e. longer than the scan itself for small buffers)
                    // warps interleaved across the CTAs, so a batch smaller than the grid still uses every SM
                    nb = ((unsigned long long)(threadIdx.x >> 5) * gridDim.x + blockIdx.x) * 32u + (threadIdx.x & 31u);
                    first = false;
                } else {
                    // later buffers: one atomic per group of lanes that arrive here together
                    const u32 act = __activemask(), ln = threadIdx.x & 31u, leader = (u32)__ffs((int)act) - 1u;
                    unsigned long long base = 0;
                    if (ln == leader) base = atomicAdd(next_buf, (unsigned long long)__popc(act));
                    base = __shfl_sync(act, base, leader);
                    nb = base + (unsigned long long)gridDim.x * blockDim.x + (u32)__popc(act & ((1u << ln) - 1u));
                }
                if (nb >= items) break;
                const u32 b = cp_.order ? cp_.order[(u32)(nb / npk)] : (u32)(nb / npk);
                o0 = cp_.off[b];
                l = (u32)(cp_.off[b + 1] - o0);
                if (l && passes_gates(db_, l)) break;
            }
            if (nb >= items) break;
            buf = cp_.order ? cp_.order[(u32)(nb / npk)] : (u32)(nb / npk);
            eng = (u32)(nb % npk);
            {
                const DevMcc &me = pk.m[eng];
                rowbytes = (WIDE16 ? 2u : 1u) << me.as;
                accept_limit = me.accept_limit_8;
                rm_base = smem + (size_t)eng * MCC_RM_BYTES;
                tab = smem + pk.tab_off[eng] + (me.tab_abs & 3u);
                tab_sh = (u32)__cvta_generic_to_shared(tab);
            }
            len = l;
            const uintptr_t ap = reinterpret_cast<uintptr_t>(cp_.text + o0);
            chunk = reinterpret_cast<const uint4 *>(ap & ~(uintptr_t)15);
            i = -(long long)(ap & 15);
            a[0] = chunk[0];             // over-reads stay inside the padded text (BACK_PAD)
            a[1] = chunk[1]; a[2] = chunk[2]; a[3] = chunk[3];
            chunk += 4;
            s = pk.m[eng].start;
            have = true;
        }
#pragma unroll
      for (int q = 0; q < 4; q++) {
        if (!have) break;  // the buffer ended in an earlier quarter of this trip
        // ---- one aligned 16-byte chunk: bytes [j0, j1) of it belong to the buffer ----
        const uint4 v = a[q];
        const long long rem = (long long)len - i;  // bytes from chunk byte 0 to the end of the buffer
        // three chunks (48 steps) of loads in flight while this one is stepped: a step is ~40 cycles of dependent
        // latency, a miss to HBM ~1000.  (A rotating FIFO -- cur = q1; q1 = q2; q2 = q3 -- makes the move wait for
        // the load issued one trip earlier: 20 % of the stall samples sat there.)
        if (rem > 64) a[q] = *chunk;
        chunk++;
        const u32 j0 = i < 0 ? (u32)(-i) : 0u;
        const u32 j1 = rem < 16 ? (u32)rem : 16u;
        const u32 valid = ((1u << j1) - 1u) & ~((1u << j0) - 1u);
        const u32 wv[4] = {v.x, v.y, v.z, v.w};
        const u32 s0 = s;
        bool exact = SHERMAN;
        if (!SHERMAN) {
            u32 seen = 0;  // 16-bit: OR of the entries taken; 8-bit: highest state entered
#pragma unroll
            for (int j = 0; j < 16; j++) {
                // byte j of the chunk, times 128 (one 32-word row of the symbol table per byte value), plus my bank
                const u32 w = wv[j >> 2], k = 8 * (j & 3);
                const u32 ra = ((k >= 7 ? w >> (k - 7) : w << (7 - k)) & 0x7f80u) | lane4;
                const u32 sym = *reinterpret_cast<const u32 *>(rm_base + ra);
                u32 e;
                if (SMEM) {
                    // row offset + (symbol offset + table address): the second term is off the dependent chain, which
                    // is then mask -> multiply-add -> shared-memory read
                    u32 addr;  // written as PTX so the compiler keeps the association
                    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(addr) : "r"(WIDE16 ? (s & mcclellan::STATE_MASK) : s), "r"(rowbytes), "r"(sym));
                    if (WIDE16) { unsigned short t; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(t) : "r"(addr)); e = t; }
                    else asm volatile("ld.shared.u8 %0, [%1];" : "=r"(e) : "r"(addr));
                } else if (WIDE16) e = *reinterpret_cast<const u16 *>(tab + (s & mcclellan::STATE_MASK) * rowbytes + sym);
                else e = tab[s * rowbytes + sym];
                s = ((valid >> j) & 1u) ? e : s;   // bytes outside the buffer freeze the state (flags of s stay as they were)
                if (WIDE16) seen |= s; else seen = max(seen, s);
            }
            if (WIDE16) { exact = (seen & mcclellan::ACCEPT_FLAG) != 0; s &= mcclellan::STATE_MASK; }
            else exact = seen >= accept_limit;
        }
        if (exact) {
            // the reference's loop, byte by byte, raising the accepts (mcclellan.c:267-289 / :482-506)
            s = s0;
            for (u32 j = j0; j < j1 && s; j++) {
                const u32 w = j < 8 ? (j < 4 ? v.x : v.y) : (j < 12 ? v.z : v.w);
                const u32 e = mcc_step<WIDE16>(db_.bc, pk.m[eng], tab, rowbytes, s,
                                               reinterpret_cast<const u32 *>(rm_base)[((w >> (8 * (j & 3))) & 0xffu) * 32 + (threadIdx.x & 31u)] - tab_sh);
                s = WIDE16 ? (e & mcclellan::STATE_MASK) : e;
                if (WIDE16 ? (e & mcclellan::ACCEPT_FLAG) != 0 : e >= accept_limit)
                    mcc_accept(evq, db_.bc, pk.m[eng], s, buf, (u32)(i + j + 1), false);
            }
        }
        i += 16;
        if (i >= (long long)len || !s) {
            // ---- end of buffer (or dead: nothing can switch on again, mcclellan.c:240-243) ----
            if (s && pk.m[eng].accepts_eod) mcc_accept(evq, db_.bc, pk.m[eng], s, buf, len, true);
            have = false;
        }
      }
    }
}
