// accel.cu -- shufti / truffle / vermicelli byte-class searches as warp-level tests (part of libhs_gpu.so).
//
// The reference's acceleration primitives (/root/reference/src/nfa) answer "where is the first (last) byte
// of [buf, end) that belongs to a class":
//   shufti     lo[c & 15] & hi[c >> 4] != 0                         shufti.cpp:44-70, shufti_simd.hpp:89-138
//   truffle    (c & 0x80 ? highset : highclear)[c & 15] bit ((c >> 4) & 7)   x86/truffle.hpp:37-62, trufflecompile.cpp:59-73
//   vermicelli (c & casemask) == ch ; double: adjacent pair, a trailing c1 counts   vermicelli_simd.cpp:82-622
// with the return convention: forward -> pointer to the hit or buf_end; reverse -> pointer to the hit or buf - 1.
// On the CPU these are one SIMD compare chain per 16..64 bytes.  Here every lane classifies 16 bytes
// (one LDG.128), the warp votes with a ballot, and the grid keeps the minimum (maximum) hit offset with one
// atomic per warp that saw a hit; blocks whose range lies behind an already-published hit exit early.
#include "hs_gpu.h"

#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>

namespace {

enum Kind { SHUFTI = 0, TRUFFLE = 1, VERM = 2, NVERM = 3, DVERM = 4, DSHUFTI = 5 };

struct ClassParams {
    uint8_t a[16];  // shufti lo / truffle highclear
    uint8_t b[16];  // shufti hi / truffle highset
    uint8_t a2[16], b2[16];  // double shufti: lo / hi nibble masks of the second byte
    uint8_t c1, c2, casemask, kind;
};

__device__ __forceinline__ bool in_class(const ClassParams &p, const uint8_t *sa, const uint8_t *sb, uint8_t c) {
    switch (p.kind) {
    case SHUFTI: return (sa[c & 15] & sb[c >> 4]) != 0;
    case TRUFFLE: return ((((c & 0x80) ? sb : sa)[c & 15] >> ((c >> 4) & 7)) & 1) != 0;
    case VERM: return (uint8_t)(c & p.casemask) == p.c1;
    default: return (uint8_t)(c & p.casemask) != p.c1;  // NVERM
    }
}

// result[0]: forward = smallest hit offset (init len), reverse = largest hit offset + 1 (init 0)
template <bool REVERSE>
__global__ void k_class_search(const uint8_t *buf, unsigned long long len, ClassParams p, unsigned long long *result) {
    __shared__ uint8_t sa[16], sb[16], sa2[16], sb2[16];
    if (threadIdx.x < 16) {
        sa[threadIdx.x] = p.a[threadIdx.x]; sb[threadIdx.x] = p.b[threadIdx.x];
        sa2[threadIdx.x] = p.a2[threadIdx.x]; sb2[threadIdx.x] = p.b2[threadIdx.x];
    }
    __syncthreads();
    const unsigned long long chunks = (len + 15) / 16;
    const unsigned long long per_block = (chunks + gridDim.x - 1) / gridDim.x;
    unsigned long long first = (unsigned long long)blockIdx.x * per_block, last = first + per_block;
    if (last > chunks) last = chunks;
    const unsigned lane = threadIdx.x & 31;
    for (unsigned long long base = first + (threadIdx.x & ~31u); base < last; base += blockDim.x) {
        // early exit: a better hit is already known
        unsigned long long best = *(volatile unsigned long long *)result;
        if (!REVERSE && best <= base * 16) return;
        unsigned long long ch = REVERSE ? (first + last - 1 - (base - first) - lane) : base + lane;
        // (reverse walks the block's range from its end; lanes keep ascending order inside the warp's 512 bytes)
        if (REVERSE) {
            const unsigned long long top = last - 1 - (base - first);  // highest chunk of this warp step
            ch = top >= lane ? top - lane : ~0ull;
            if (best > (top + 1) * 16) return;
        }
        unsigned hits = 0;
        if (ch != ~0ull && ch >= first && ch < last) {
            const unsigned long long off = ch * 16;
            uint4 v = *reinterpret_cast<const uint4 *>(buf + off);  // buffer is padded to a multiple of 16 + 16
            const unsigned w[5] = {v.x, v.y, v.z, v.w, *reinterpret_cast<const unsigned *>(buf + off + 16)};
#pragma unroll
            for (int i = 0; i < 16; i++) {
                if (off + i >= len) break;
                const uint8_t c = (uint8_t)(w[i >> 2] >> (8 * (i & 3)));
                bool hit;
                if (p.kind == DVERM) {
                    const uint8_t n = (uint8_t)(w[(i + 1) >> 2] >> (8 * ((i + 1) & 3)));
                    const bool is1 = (uint8_t)(c & p.casemask) == p.c1;
                    hit = is1 && (off + i + 1 >= len || (uint8_t)(n & p.casemask) == p.c2);
                } else if (p.kind == DSHUFTI) {
                    // blockDoubleMask (src/nfa/x86/shufti.hpp:49-74): the masks hold CLEAR bits for members; a pair
                    // bucket k matches at i when bit k is clear for byte i under mask 1 and for byte i + 1 under mask 2;
                    // past the end of the data the second byte admits everything (a trailing first byte counts)
                    const uint8_t n = (uint8_t)(w[(i + 1) >> 2] >> (8 * ((i + 1) & 3)));
                    const uint8_t m1 = (uint8_t)(sa[c & 15] | sb[c >> 4]);
                    const uint8_t m2 = off + i + 1 >= len ? (uint8_t)0 : (uint8_t)(sa2[n & 15] | sb2[n >> 4]);
                    hit = (uint8_t)(m1 | m2) != 0xff;
                } else {
                    hit = in_class(p, sa, sb, c);
                }
                hits |= (hit ? 1u : 0u) << i;
            }
        }
        const unsigned vote = __ballot_sync(0xffffffffu, hits != 0);
        if (vote) {
            if (!REVERSE) {
                const unsigned src = __ffs(vote) - 1;  // lowest chunk with a hit
                const unsigned h = __shfl_sync(0xffffffffu, hits, src);
                if (lane == 0) atomicMin(result, (base + src) * 16 + (__ffs(h) - 1));
            } else {
                const unsigned src = __ffs(vote) - 1;  // lane 0 holds the highest chunk
                const unsigned h = __shfl_sync(0xffffffffu, hits, src);
                const unsigned long long chs = __shfl_sync(0xffffffffu, ch, src);
                if (lane == 0) atomicMax(result, chs * 16 + (31 - __clz(h)) + 1);
            }
            return;  // every later chunk of this warp is worse; other warps check `best`
        }
    }
}


// ---- k_class_scan: the single-byte classes (shufti, truffle, vermicelli, negated vermicelli) and the double vermicelli ----
// Every one of them is "byte c is in a set" (the double form: c in set 1 and the NEXT byte in set 2), so the class is
// turned on the host into a 256-entry table -- bit 0 = member of set 1, bit 16 = member of set 2 -- and the kernel is
// one conflict-free shared-memory read per byte: the table is replicated once per bank (word c * 32 + lane, 32 KiB), as
// k_mcclellan's symbol table.  A warp step is 2 KiB: four coalesced LDG.128 per lane issued together, then 4 instructions
// per byte -- PRMT (byte), IMAD (row address), LDS, IMAD (mask * 2 + flag) -- one on the ALU pipe, two on the FMA pipe.
// Forward: the grid publishes the smallest hit offset (atomicMin); blocks own contiguous ranges and look at the published
// value every few steps to stop early.  Reverse: the largest hit offset + 1 (atomicMax), ranges walked from their end.
constexpr int CS_THREADS = 256;
constexpr int CS_TILE_CHUNKS = 128;  // 16-byte chunks per warp step (4 per lane)

template <bool REVERSE, int MODE>
__global__ void __launch_bounds__(CS_THREADS) k_class_scan(const uint8_t *buf, unsigned long long len, const unsigned *table256,
                                                           unsigned long long *result) {
    __shared__ unsigned cs_tab[256 * 32];  // [256][32]; static, so that its address is an immediate of the LDS
    for (unsigned i = threadIdx.x; i < 256 * 32; i += blockDim.x) cs_tab[i] = table256[i >> 5];
    __syncthreads();
    constexpr bool DOUBLE = MODE != 0;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
    // my bank's copy: the entry of byte c sits at shared address tab_l + c * 128
    const unsigned tab_l = (unsigned)__cvta_generic_to_shared(cs_tab) + lane * 4;
    const unsigned long long chunks = (len + 15) / 16;
    const unsigned long long tiles = (chunks + CS_TILE_CHUNKS - 1) / CS_TILE_CHUNKS;
    const unsigned long long per_block = (tiles + gridDim.x - 1) / gridDim.x;
    const unsigned long long t0 = (unsigned long long)blockIdx.x * per_block;
    unsigned long long t1 = t0 + per_block;
    if (t1 > tiles) t1 = tiles;
    unsigned step = 0;
    for (unsigned long long it = warp; t0 + it < t1; it += warps, step++) {
        const unsigned long long tile = REVERSE ? t1 - 1 - it : t0 + it;
        if ((step & 3) == 0) {  // a better hit is already known: nothing in this warp's remaining tiles can beat it
            const unsigned long long best = *(volatile unsigned long long *)result;
            if (!REVERSE && best <= tile * (CS_TILE_CHUNKS * 16ull)) return;
            if (REVERSE && best >= (tile + 1) * (CS_TILE_CHUNKS * 16ull)) return;
        }
        // a tile that ends before the last chunk of the data needs no bounds at all (the common case)
        const bool inner = (tile + 1) * CS_TILE_CHUNKS < chunks;
        const unsigned char *tp = buf + tile * (CS_TILE_CHUNKS * 16ull) + lane * 16u;
        uint4 v[4];
        unsigned nxt[4];
        if (inner) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                v[k] = *reinterpret_cast<const uint4 *>(tp + k * 512);
                if (DOUBLE) nxt[k] = tp[k * 512 + 16];
            }
        } else {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const unsigned long long ch = tile * CS_TILE_CHUNKS + (unsigned)k * 32 + lane;
                v[k] = ch < chunks ? *reinterpret_cast<const uint4 *>(tp + k * 512) : make_uint4(0, 0, 0, 0);
                if (DOUBLE) nxt[k] = ch < chunks ? tp[k * 512 + 16] : 0u;  // the staged data is padded: always readable
            }
        }
        unsigned hits[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const unsigned w[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
            unsigned a1 = 0, a2 = 0;  // bit 15 - i = byte i of the chunk
            unsigned h;
            if (MODE != 2) {
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    // byte i by one PRMT (ALU pipe); its row and the mask update by multiply-adds (FMA pipe); the flag of an
                    // entry is its bit 0 (set 1) and bit 16 (set 2), so that mask * 2 + entry appends both at once
                    const unsigned c = __byte_perm(w[i >> 2], 0, 0x4440 + (i & 3));
                    unsigned e;
                    asm volatile("{ .reg .u32 a; mad.lo.u32 a, %1, 128, %2; ld.shared.u32 %0, [a]; }" : "=r"(e) : "r"(c), "r"(tab_l));
                    if (!DOUBLE) a1 = a1 * 2u + e;
                    else { a1 = a1 * 2u + (e & 1u); a2 = a2 * 2u + (e >> 16); }
                }
                h = a1;
                if (DOUBLE) {
                    const unsigned e16 = cs_tab[nxt[k] * 32 + lane];            // the byte after the chunk
                    const unsigned need = ((a2 << 1) | (e16 >> 16)) & 0xffffu;   // set 2 holds the byte after byte i
                    h = a1 & need;
                }
            } else {
                // double shufti (blockDoubleMask, src/nfa/x86/shufti.hpp:49-74): entry = mask-1 byte | mask-2 byte << 8, both
                // with CLEAR bits for members; byte i hits when (m1[c_i] | m2[c_i+1]) leaves some pair bucket clear
                unsigned e[17];
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    const unsigned c = __byte_perm(w[i >> 2], 0, 0x4440 + (i & 3));
                    asm volatile("{ .reg .u32 a; mad.lo.u32 a, %1, 128, %2; ld.shared.u32 %0, [a]; }" : "=r"(e[i]) : "r"(c), "r"(tab_l));
                }
                e[16] = cs_tab[nxt[k] * 32 + lane];
                h = 0;
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    const unsigned pr = (e[i] | (e[i + 1] >> 8)) & 0xffu;     // 0xff = no bucket admits the pair
                    h = h * 2u + (((pr + 1u) >> 8) ^ 1u);
                }
                a1 = 0;
                if (!inner) {
#pragma unroll
                    for (int i = 0; i < 16; i++) a1 = a1 * 2u + ((((e[i] & 0xffu) + 1u) >> 8) ^ 1u);  // first byte alone (last-byte rule)
                }
            }
            if (!inner) {
                const unsigned long long ch = tile * CS_TILE_CHUNKS + (unsigned)k * 32 + lane;
                // the last byte of the data has no successor: a member of set 1 there counts (vermicelliDoubleExec)
                if (DOUBLE && len - 1 >= ch * 16 && len - 1 < ch * 16 + 16) h |= a1 & (0x8000u >> (unsigned)(len - 1 - ch * 16));
                // bytes past the end of the data (zero padding) are not part of the search
                if (ch >= chunks) h = 0;
                else if (ch * 16 + 16 > len) h &= 0xffffu << (unsigned)(ch * 16 + 16 - len);
            }
            hits[k] = h & 0xffffu;
        }
        if (!REVERSE) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const unsigned vote = __ballot_sync(0xffffffffu, hits[k] != 0);
                if (vote) {
                    const unsigned src = __ffs(vote) - 1;
                    const unsigned h = __shfl_sync(0xffffffffu, hits[k], src);
                    if (lane == 0) atomicMin(result, (tile * CS_TILE_CHUNKS + (unsigned)k * 32 + src) * 16ull + (unsigned)(__clz(h) - 16));
                    return;  // this warp's later tiles lie behind this hit
                }
            }
        } else {
#pragma unroll
            for (int k = 3; k >= 0; k--) {
                const unsigned vote = __ballot_sync(0xffffffffu, hits[k] != 0);
                if (vote) {
                    const unsigned src = 31 - __clz(vote);
                    const unsigned h = __shfl_sync(0xffffffffu, hits[k], src);
                    if (lane == 0) atomicMax(result, (tile * CS_TILE_CHUNKS + (unsigned)k * 32 + src) * 16ull + (unsigned)(15 - (__ffs(h) - 1)) + 1ull);
                    return;
                }
            }
        }
    }
}

struct Workspace {
    uint8_t *d_buf = nullptr;
    size_t cap = 0;
    unsigned long long *d_res = nullptr;
    unsigned *d_tab = nullptr;  // 256-entry class table of k_class_scan
    size_t len = 0;
};

Workspace &workspace() {
    static thread_local Workspace ws;
    return ws;
}

// stage the data on the device (padded to a multiple of 16 + 32 zero bytes)
bool stage(const uint8_t *buf, size_t len) {
    Workspace &ws = workspace();
    const size_t padded = (len + 15) / 16 * 16 + 32;
    if (padded > ws.cap) {
        if (ws.d_buf) cudaFree(ws.d_buf);
        if (cudaMalloc((void **)&ws.d_buf, padded + padded / 4) != cudaSuccess) { ws.d_buf = nullptr; ws.cap = 0; return false; }
        ws.cap = padded + padded / 4;
    }
    if (!ws.d_res && cudaMalloc((void **)&ws.d_res, 8) != cudaSuccess) return false;
    if (cudaMemcpy(ws.d_buf, buf, len, cudaMemcpyHostToDevice) != cudaSuccess) return false;
    cudaMemset(ws.d_buf + len, 0, padded - len);
    ws.len = len;
    return true;
}

// one search over the staged data; `repeats` > 1 re-runs it and reports the mean kernel time (CUDA events)
long long search_staged(const ClassParams &p, bool reverse, int repeats, float *ms) {
    Workspace &ws = workspace();
    const size_t len = ws.len;
    if (!len) return reverse ? -1 : 0;
    const unsigned long long chunks = (len + 15) / 16;
    unsigned grid = (unsigned)((chunks + 255) / 256);
    if (grid > 148 * 8) grid = 148 * 8;
    // every kind but the double shufti is a per-byte set test: one table, one kernel (k_class_scan)
    const bool table_kind = !(p.kind == DSHUFTI && reverse);
    const bool dbl = p.kind == DVERM;
    const bool dsh = p.kind == DSHUFTI;
    if (table_kind) {
        unsigned tab[256];
        for (unsigned c = 0; c < 256; c++) {
            bool in1, in2 = false;
            switch (p.kind) {
            case SHUFTI: in1 = (p.a[c & 15] & p.b[c >> 4]) != 0; break;
            case TRUFFLE: in1 = ((((c & 0x80) ? p.b : p.a)[c & 15] >> ((c >> 4) & 7)) & 1) != 0; break;
            case VERM: in1 = (uint8_t)(c & p.casemask) == p.c1; break;
            case NVERM: in1 = (uint8_t)(c & p.casemask) != p.c1; break;
            case DSHUFTI: in1 = false; break;
            default: in1 = (uint8_t)(c & p.casemask) == p.c1; in2 = (uint8_t)(c & p.casemask) == p.c2; break;
            }
            tab[c] = (in1 ? 1u : 0u) | (in2 ? 0x10000u : 0u);
            if (dsh) tab[c] = (unsigned)(uint8_t)(p.a[c & 15] | p.b[c >> 4]) | ((unsigned)(uint8_t)(p.a2[c & 15] | p.b2[c >> 4]) << 8);
        }
        if (!ws.d_tab && cudaMalloc((void **)&ws.d_tab, sizeof(tab)) != cudaSuccess) return -2;
        cudaMemcpy(ws.d_tab, tab, sizeof(tab), cudaMemcpyHostToDevice);
        const unsigned long long tiles = (chunks + CS_TILE_CHUNKS - 1) / CS_TILE_CHUNKS;
        const unsigned long long want = (tiles + (CS_THREADS / 32) * 4 - 1) / ((CS_THREADS / 32) * 4);  // >= 4 steps per warp
        grid = (unsigned)(want < 1 ? 1 : want > 148 * 6 ? 148 * 6 : want);
    }
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ms) { cudaEventCreate(&e0); cudaEventCreate(&e1); }
    float total = 0;
    unsigned long long res = 0;
    for (int r = 0; r < (repeats > 0 ? repeats : 1); r++) {
        const unsigned long long init = reverse ? 0ull : (unsigned long long)len;
        cudaMemcpy(ws.d_res, &init, 8, cudaMemcpyHostToDevice);
        if (ms) cudaEventRecord(e0);
        if (table_kind) {
            if (dsh) k_class_scan<false, 2><<<grid, CS_THREADS>>>(ws.d_buf, len, ws.d_tab, ws.d_res);
            else if (dbl) k_class_scan<false, 1><<<grid, CS_THREADS>>>(ws.d_buf, len, ws.d_tab, ws.d_res);
            else if (reverse) k_class_scan<true, 0><<<grid, CS_THREADS>>>(ws.d_buf, len, ws.d_tab, ws.d_res);
            else k_class_scan<false, 0><<<grid, CS_THREADS>>>(ws.d_buf, len, ws.d_tab, ws.d_res);
        } else if (reverse) k_class_search<true><<<grid, 256>>>(ws.d_buf, len, p, ws.d_res);
        else k_class_search<false><<<grid, 256>>>(ws.d_buf, len, p, ws.d_res);
        if (ms) {
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float t = 0;
            cudaEventElapsedTime(&t, e0, e1);
            total += t;
        }
    }
    if (cudaMemcpy(&res, ws.d_res, 8, cudaMemcpyDeviceToHost) != cudaSuccess) return -2;
    if (ms) { *ms = total / (float)(repeats > 0 ? repeats : 1); cudaEventDestroy(e0); cudaEventDestroy(e1); }
    return reverse ? (long long)res - 1 : (long long)res;
}

long long run_search(const ClassParams &p, const uint8_t *buf, size_t len, bool reverse) {
    if (!buf && len) return -2;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return -2; }
    if (!len) return reverse ? -1 : 0;
    if (!stage(buf, len)) return -2;
    return search_staged(p, reverse, 1, nullptr);
}

ClassParams make_params(int kind, const uint8_t *m1, const uint8_t *m2, const uint8_t *m3, const uint8_t *m4, uint8_t c1, uint8_t c2, int nocase) {
    ClassParams p{};
    if (m1) std::memcpy(p.a, m1, 16);
    if (m2) std::memcpy(p.b, m2, 16);
    if (m3) std::memcpy(p.a2, m3, 16);
    if (m4) std::memcpy(p.b2, m4, 16);
    p.casemask = nocase ? 0xdf : 0xff;
    p.c1 = (uint8_t)(c1 & p.casemask);
    p.c2 = (uint8_t)(c2 & p.casemask);
    p.kind = (uint8_t)kind;
    return p;
}

// ---- hs_gpu_shufti_noalloc: the same search with NO device allocation (no counterpart in the reference) ---------------
// Everything hs_gpu_shufti keeps in its cudaMalloc'ed workspace -- the result cell, the 256-entry class table, the staged
// and padded copy of the data -- lives in ONE file-scope device array of fixed size, so a caller that may not allocate
// (under CUDA graph capture, inside a serving loop, in a harness that forbids cudaMalloc) can still search.  Every GPU
// that loads libhs_gpu.so holds the array until the process exits.  In host code `g_noalloc_scratch` is only the
// host-side stand-in of the symbol: its device address comes from cudaGetSymbolAddress, once per device, with that device
// current, and that address is what the kernel is given.
constexpr size_t NA_RES_OFF = 0, NA_TAB_OFF = 256, NA_TEXT_OFF = 256 + 1024;
constexpr size_t NA_PAD = 48;  // data padded to a multiple of 16 + 32 zero bytes, as stage() does
constexpr size_t NA_BYTES = NA_TEXT_OFF + HS_GPU_NOALLOC_MAX_BYTES + NA_PAD;
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_noalloc_mutex;                       // one search at a time per process: every call shares the array
unsigned char *g_noalloc_ptr[NA_MAX_DEVICES];     // device address of g_noalloc_scratch on device d (nullptr: not asked yet)

// device address of the scratch array on the CURRENT device (asked once per device)
int noalloc_scratch(unsigned char **out) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= NA_MAX_DEVICES) { cudaGetLastError(); return HS_ARCH_ERROR; }
    if (!g_noalloc_ptr[dev]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            std::fprintf(stderr, "hs_gpu_shufti_noalloc: cudaGetSymbolAddress failed on device %d: %s\n", dev, cudaGetErrorString(e));
            cudaGetLastError();
            return HS_ARCH_ERROR;
        }
        g_noalloc_ptr[dev] = static_cast<unsigned char *>(p);
    }
    *out = g_noalloc_ptr[dev];
    return HS_SUCCESS;
}

}  // namespace

extern "C" {

long long hs_gpu_shufti(const uint8_t lo[16], const uint8_t hi[16], const uint8_t *buf, size_t len, int reverse) {
    ClassParams p{};
    std::memcpy(p.a, lo, 16);
    std::memcpy(p.b, hi, 16);
    p.kind = SHUFTI;
    return run_search(p, buf, len, reverse != 0);
}

long long hs_gpu_truffle(const uint8_t mask1[16], const uint8_t mask2[16], const uint8_t *buf, size_t len, int reverse) {
    ClassParams p{};
    std::memcpy(p.a, mask1, 16);
    std::memcpy(p.b, mask2, 16);
    p.kind = TRUFFLE;
    return run_search(p, buf, len, reverse != 0);
}

long long hs_gpu_vermicelli(uint8_t c, int nocase, const uint8_t *buf, size_t len, int reverse, int negate) {
    ClassParams p{};
    p.casemask = nocase ? 0xdf : 0xff;
    p.c1 = (uint8_t)(c & p.casemask);
    p.kind = negate ? NVERM : VERM;
    return run_search(p, buf, len, reverse != 0);
}

long long hs_gpu_vermicelli_double(uint8_t c1, uint8_t c2, int nocase, const uint8_t *buf, size_t len) {
    ClassParams p{};
    p.casemask = nocase ? 0xdf : 0xff;
    p.c1 = (uint8_t)(c1 & p.casemask);
    p.c2 = (uint8_t)(c2 & p.casemask);
    p.kind = DVERM;
    return run_search(p, buf, len, false);
}

long long hs_gpu_shufti_double(const uint8_t lo1[16], const uint8_t hi1[16], const uint8_t lo2[16], const uint8_t hi2[16],
                               const uint8_t *buf, size_t len) {
    return run_search(make_params(DSHUFTI, lo1, hi1, lo2, hi2, 0, 0, 0), buf, len, false);
}

// Measurement entry points (tools/accel_bench.py): the data is staged once, then searched `repeats` times with CUDA
// events around each kernel.  kind: 0 shufti, 1 truffle, 2 vermicelli, 3 negated vermicelli, 4 double vermicelli,
// 5 double shufti; masks = up to four 16-byte masks back to back.
int hs_gpu_accel_stage(const uint8_t *buf, size_t len) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return -11; }
    return stage(buf, len) ? 0 : -2;
}
long long hs_gpu_accel_search_staged(int kind, const uint8_t *masks, uint8_t c1, uint8_t c2, int nocase, int reverse,
                                     int repeats, float *kernel_ms) {
    if (kind < 0 || kind > 5) return -2;
    const ClassParams p = make_params(kind, masks, masks ? masks + 16 : nullptr, masks ? masks + 32 : nullptr,
                                      masks ? masks + 48 : nullptr, c1, c2, nocase);
    return search_staged(p, reverse != 0, repeats, kernel_ms);
}

// shuftiExec / rshuftiExec (src/nfa/shufti.h) with no device allocation: see the comment at g_noalloc_scratch.  `buf` is
// host memory, *result receives what hs_gpu_shufti returns.  Calls are serialised inside the process and complete before
// they return; the array keeps the last call's contents, so the result cell, the table and the padding behind the data
// are rewritten by every call.
hs_error_t hs_gpu_shufti_noalloc(const uint8_t lo[16], const uint8_t hi[16], const uint8_t *buf, size_t len, int reverse,
                                 long long *result) {
    if (!lo || !hi || !result || (!buf && len)) return HS_INVALID;
    if (len > HS_GPU_NOALLOC_MAX_BYTES) {
        std::fprintf(stderr, "hs_gpu_shufti_noalloc: %zu bytes exceed the fixed scratch of %u bytes (HS_GPU_NOALLOC_MAX_BYTES)\n",
                     len, (unsigned)HS_GPU_NOALLOC_MAX_BYTES);
        return HS_INSUFFICIENT_SPACE;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return HS_ARCH_ERROR; }
    if (!len) { *result = reverse ? -1 : 0; return HS_SUCCESS; }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    unsigned char *base = nullptr;
    const int rv = noalloc_scratch(&base);
    if (rv != HS_SUCCESS) return rv;
    unsigned long long *d_res = reinterpret_cast<unsigned long long *>(base + NA_RES_OFF);
    unsigned *d_tab = reinterpret_cast<unsigned *>(base + NA_TAB_OFF);
    unsigned char *d_text = base + NA_TEXT_OFF;
    unsigned tab[256];
    for (unsigned c = 0; c < 256; c++) tab[c] = (lo[c & 15] & hi[c >> 4]) != 0 ? 1u : 0u;
    const size_t padded = (len + 15) / 16 * 16 + 32;
    const unsigned long long init = reverse ? 0ull : (unsigned long long)len;
    if (cudaMemcpy(d_res, &init, 8, cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemcpy(d_tab, tab, sizeof(tab), cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemcpy(d_text, buf, len, cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemset(d_text + len, 0, padded - len) != cudaSuccess) {
        cudaGetLastError();
        return HS_UNKNOWN_ERROR;
    }
    const unsigned long long chunks = (len + 15) / 16;
    const unsigned long long tiles = (chunks + CS_TILE_CHUNKS - 1) / CS_TILE_CHUNKS;
    const unsigned long long want = (tiles + (CS_THREADS / 32) * 4 - 1) / ((CS_THREADS / 32) * 4);
    const unsigned grid = (unsigned)(want < 1 ? 1 : want > 148 * 6 ? 148 * 6 : want);
    if (reverse) k_class_scan<true, 0><<<grid, CS_THREADS>>>(d_text, len, d_tab, d_res);
    else k_class_scan<false, 0><<<grid, CS_THREADS>>>(d_text, len, d_tab, d_res);
    unsigned long long res = 0;
    if (cudaGetLastError() != cudaSuccess || cudaMemcpy(&res, d_res, 8, cudaMemcpyDeviceToHost) != cudaSuccess) {
        cudaGetLastError();
        return HS_UNKNOWN_ERROR;
    }
    *result = reverse ? (long long)res - 1 : (long long)res;
    return HS_SUCCESS;
}

}  // extern "C"
