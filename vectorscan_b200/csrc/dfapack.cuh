// dfapack.cuh -- several McClellan DFAs stepped TOGETHER over one buffer per thread (included by hs_gpu.cu behind
// mcclellan.cuh; used for the outfix engines of full-Rose databases, launch_outfixes).
//
// Why: one DFA step is a dependent shared-memory read, so a thread that steps ONE engine (k_mcclellan) waits out the
// read's latency on every byte, and a multi-engine database pays that once per engine and per launch.  Here a thread
// carries D states (D = 2, 4 or 8 engines of the same database) through its buffer: the byte is fetched and classified
// once, the D table reads are independent of each other (D chains in flight per thread), and a step costs 5
// instructions instead of 8.
//
// What is computed is exactly what k_mcclellan computes per engine (mcclellan.cuh, src/nfa/mcclellan.c:227-366,445-565):
//     s_d = start_anchored_d;  per byte: e = succ_d[s_d][class(byte)];  accept flag -> reports of the new state at i + 1
// on tables RE-LAID AT UPLOAD (host: build_dfa_packs in hs_gpu.cu) from the engines' own bytecode:
//   * one JOINT alphabet per pack: two bytes share a joint class when every engine of the pack maps them to the same
//     symbol (remap[]); the class of a byte comes from ONE bank-replicated table (conflict-free, as k_mcclellan's symbol
//     table) holding the class's byte offset inside a row;
//   * every engine's successor table on that alphabet, 16-bit entries `next state | 0x8000 if next accepts` (8-bit
//     engines: next >= accept_limit_8), rows of 2^row_shift bytes, all engines' tables side by side in shared memory.
// A chunk of 16 bytes is stepped branch-free for all D engines while OR-ing the entries; an engine that passed an
// accept in the chunk (rare) is stepped again from its saved state by the exact byte loop, which raises the events
// {buffer, to, report list} for k_limex_report -- the scheme of k_mcclellan.  Engines with Sherman states, wide states,
// EOD accepts or tables too large for shared memory are not packed (they keep k_mcclellan).
//
// MEASURED (C3, 2 000 regexes as 114 engines, 256 MiB, B200): 18.9 ms against 18.5 ms for k_mcclellan_pack -- no gain: the
// joint alphabet doubles the row size of most 16-bit engines, so only the small engines pack at all, and with buffers
// handed out by length the single-chain kernel already keeps its lanes busy.  Option "pack" = 2 selects this kernel; the
// default is 1 (k_mcclellan_pack).  ncu (gpurun_out r2q): 5.2 shared-memory wavefronts per warp-step, issue 23 % at D = 4.
#pragma once

constexpr int DP_THREADS = 1024;
constexpr int DP_MAX = 8;

struct DevDfaPack {
    const u8 *blob;          // device: u32 class_off[256] (byte offset of the byte's class inside a row), then the tables
    u32 tab_off[DP_MAX];     // byte offset of engine d's table inside the blob's table area
    u32 start[DP_MAX];       // start_anchored of engine d
    DevMcc m[DP_MAX];        // the engines themselves (report lists, aux) for the accept path
    u32 n;                   // engines in the pack (<= D of the launch; the rest are padding: start 0 = dead)
    u32 row_shift;           // log2(bytes per table row)
    u32 tab_bytes;           // all tables together
};

__host__ __device__ inline size_t dfapack_smem_bytes(const DevDfaPack &p) { return MCC_RM_BYTES + (size_t)p.tab_bytes + 16; }

template <int D>
__global__ void __launch_bounds__(DP_THREADS, 1) k_dfa_pack(const DevDb db_, const DevDfaPack pk, const DevCorpus cp_,
                                                            const DevLxEvents evq, unsigned long long *next_buf) {
    extern __shared__ __align__(16) u8 smem[];
    u32 *rm_w = reinterpret_cast<u32 *>(smem);
    {
        const u32 *cls = reinterpret_cast<const u32 *>(pk.blob);
        for (u32 i = threadIdx.x; i < 256 * 32; i += blockDim.x) rm_w[i] = cls[i >> 5];
        const u32 words = (pk.tab_bytes + 3) / 4;
        const u32 *src = reinterpret_cast<const u32 *>(pk.blob + 1024);
        u32 *dst = reinterpret_cast<u32 *>(smem + MCC_RM_BYTES);
        for (u32 i = threadIdx.x; i < words; i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const u32 count = cp_.count, rs = pk.row_shift;
    const u32 lane4 = (threadIdx.x & 31u) * 4u;
    u32 tb[D];  // shared-memory address of engine d's table
#pragma unroll
    for (int d = 0; d < D; d++) tb[d] = (u32)__cvta_generic_to_shared(smem + MCC_RM_BYTES) + pk.tab_off[d];

    u32 s[D];
#pragma unroll
    for (int d = 0; d < D; d++) s[d] = 0;
    u32 buf = 0, len = 0;
    long long i = 0;
    const uint4 *chunk = nullptr;
    uint4 a[4];
    a[0] = a[1] = a[2] = a[3] = make_uint4(0, 0, 0, 0);
    bool have = false, first = true;

    for (;;) {
        if (!have) {
            // ---- next buffer of this thread: the scheme of k_mcclellan (first by thread index, then a counter) ----
            unsigned long long nb;
            u32 l = 0;
            u64 o0 = 0;
            for (;;) {
                if (first) {
                    nb = ((unsigned long long)(threadIdx.x >> 5) * gridDim.x + blockIdx.x) * 32u + (threadIdx.x & 31u);
                    first = false;
                } else {
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
                if (l && passes_gates(db_, l)) break;
            }
            if (nb >= count) break;
            buf = cp_.order ? cp_.order[nb] : (u32)nb;
            len = l;
            const uintptr_t ap = reinterpret_cast<uintptr_t>(cp_.text + o0);
            chunk = reinterpret_cast<const uint4 *>(ap & ~(uintptr_t)15);
            i = -(long long)(ap & 15);
            a[0] = chunk[0];
            a[1] = chunk[1]; a[2] = chunk[2]; a[3] = chunk[3];
            chunk += 4;
#pragma unroll
            for (int d = 0; d < D; d++) s[d] = pk.start[d];
            have = true;
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
            if (!have) break;
            const uint4 v = a[q];
            const long long rem = (long long)len - i;
            if (rem > 64) a[q] = *chunk;
            chunk++;
            const u32 j0 = i < 0 ? (u32)(-i) : 0u;
            const u32 j1 = rem < 16 ? (u32)rem : 16u;
            const u32 wv[4] = {v.x, v.y, v.z, v.w};
            u32 s0[D], seen[D];
#pragma unroll
            for (int d = 0; d < D; d++) { s0[d] = s[d]; seen[d] = 0; }
            // ---- 16 branch-free steps: classify once, D independent table reads; bytes outside the buffer freeze the
            // states (one path for every chunk: lanes of a warp sit at different places of different buffers, and a
            // separate path for boundary chunks would make the warp run both almost every time) ----
            const u32 valid = ((1u << j1) - 1u) & ~((1u << j0) - 1u);
#pragma unroll
            for (int j = 0; j < 16; j++) {
                const u32 w = wv[j >> 2], k = 8 * (j & 3);
                const u32 ra = ((k >= 7 ? w >> (k - 7) : w << (7 - k)) & 0x7f80u) | lane4;
                const u32 co = *reinterpret_cast<const u32 *>(smem + ra);   // class offset inside a row
                const bool in = (valid >> j) & 1u;
#pragma unroll
                for (int d = 0; d < D; d++) {
                    const u32 addr = ((s[d] & mcclellan::STATE_MASK) << rs) + (tb[d] + co);
                    unsigned short t;
                    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(t) : "r"(addr));
                    s[d] = in ? (u32)t : s[d];
                    seen[d] |= s[d];
                }
            }
            // ---- an engine that passed an accept: its chunk again, byte by byte, raising the events ----
#pragma unroll
            for (int d = 0; d < D; d++) {
                if (!(seen[d] & mcclellan::ACCEPT_FLAG)) continue;
                u32 st = s0[d] & mcclellan::STATE_MASK;
                for (u32 j = j0; j < j1; j++) {
                    const u32 w = j < 8 ? (j < 4 ? v.x : v.y) : (j < 12 ? v.z : v.w);
                    const u32 co = rm_w[((w >> (8 * (j & 3))) & 0xffu) * 32 + (threadIdx.x & 31u)];
                    const u32 e = *reinterpret_cast<const u16 *>(smem + MCC_RM_BYTES + pk.tab_off[d] + (st << rs) + co);
                    st = e & mcclellan::STATE_MASK;
                    if (e & mcclellan::ACCEPT_FLAG) mcc_accept(evq, db_.bc, pk.m[d], st, buf, (u32)(i + j + 1), false);
                }
            }
            i += 16;
            u32 alive = 0;
#pragma unroll
            for (int d = 0; d < D; d++) { s[d] &= mcclellan::STATE_MASK; alive |= s[d]; }   // no flag survives the chunk
            if (i >= (long long)len || !alive) have = false;   // end of the buffer, or every engine dead (mcclellan.c:240-243)
        }
    }
}
