// turbo_tables.h -- upload-time re-encoding of a database's literal matcher for the "turbo" front end.
//
// the FDR/Teddy front end is only a FILTER (it may over-approximate); exactness is decided
// by the confirm step ((conf_key & msk) == v, size/history check, groups: src/fdr/fdr_confirm_runtime.h:43-102)
// and the literal's Rose program.  So the GPU may use any SOUND filter derived from the database's own
// LitInfo records.  Two structures are derived here from them (nothing else is consulted):
//
//  1. PAIR TABLE (shared memory, 2^H1 x 8 bytes).  Let K = min(4, shortest LitInfo.size) and call the last K
//     bytes of a literal its tail.  Two neighbouring end positions e-1 and e share the K-1 text bytes
//     S = text[e-K+1 .. e-1].  Row hash(S) holds two 32-bit sets:  A, indexed by the low 5 bits of
//     text[e-K] (bit set = some tail is text[e-K] + S, i.e. a literal may end at e-1) and B, indexed by the
//     low 5 bits of text[e] (some tail is S + text[e]: a literal may end at e).  One 8-byte lookup therefore
//     filters two end positions.  Caseless/masked tail bytes are expanded into all byte values they admit.
//  2. SLOT TABLE (global memory / L2, open addressing, 32-byte slots).  Keyed by the exact K-byte tail; each
//     slot carries the LitInfo's v / msk / size and the literal's Rose program pre-decoded when it is one of
//     the three shapes the compiler emits for plain reports (REPORT | FINAL_REPORT, optionally guarded by one
//     CHECK_MED_LIT[_NOCASE]); anything else keeps its program offset and is interpreted.
//  3. TAIL TABLE (shared memory, 2^(BL-5) words of 32 bits).  Second filter tier, consulted only for the
//     positions the pair table lets through, ONE lookup per position: the word is chosen by a hash of the K-byte
//     tail.  A literal of exactly K bytes -- or one whose longer tail cannot be enumerated -- sets the bit
//     picked by the low 5 bits of that hash; a literal of more than K bytes sets the bit picked by the low 5
//     bits of the byte BEFORE its K-tail.  A position survives when either of its two bits is set.  The pair table alone passes every position whose K-tail
//     LOOKS like some literal's (it indexes the extra byte by 5 bits only); this cuts that to (nearly) the
//     positions where a literal's (K+1)-byte tail is present, so the L2-resident slot table is probed an order
//     of magnitude less often.
#pragma once
#include "hsb_dbview.h"

#include <algorithm>
#include <cstdlib>
#include <vector>

namespace hsb {
namespace turbo {

constexpr u32 PAIR_MULT = 0x9E3779B1u, SLOT_MULT = 0x85EBCA6Bu;
enum SlotKind : u32 { SLOT_EMPTY = 0, SLOT_REPORT = 1, SLOT_LIT_REPORT = 2, SLOT_LIT_NOCASE_REPORT = 3, SLOT_PROGRAM = 4 };

struct Slot {     // 32 bytes
    u64 v, msk;   // LitInfo.v / LitInfo.msk (8-byte window ending at the candidate, last byte on top)
    u32 key;      // the K tail bytes this slot was filed under (first tail byte lowest)
    u32 target;   // report id (SLOT_*REPORT) or Rose program offset (SLOT_PROGRAM)
    u32 lit_off;  // bytecode offset of the full literal text (SLOT_LIT_*)
    u32 meta;     // size | kind << 8 | lit_len << 16
};
static_assert(sizeof(Slot) == 32, "slot layout");

// K-1 shared bytes packed the way the kernel's PRMT selectors pack them
HSB_HD u32 pack_shared(u32 K, u32 s0, u32 s1, u32 s2) {
    switch (K) {
    case 4: return s1 | (s0 << 8) | (s2 << 16) | (s2 << 24);
    case 3: return s0 | (s1 << 8) | (s0 << 16) | (s1 << 24);
    case 2: return s0 * 0x01010101u;
    default: return 0;
    }
}
// Row of the pair table.  K == 3 (two shared bytes): the low 7 bits of each byte index a 2^14-row table
// directly -- no collisions at all on ASCII text, 4-to-1 aliasing on binary data (what a hash would give too).
// Other K: multiplicative hash of the packed shared bytes.
HSB_HD u32 pair_row(u32 K, u32 packed, u32 H1) {
    if (K == 3) return (packed & 0x7f) | ((packed >> 1) & 0x3f80);
    return (packed * PAIR_MULT) >> (32 - H1);
}
HSB_HD u32 slot_home(u32 key, u32 H2) { return (key * SLOT_MULT) >> (32 - H2); }
// tail table addressing: h = tail_hash(K-tail, first tail byte lowest); the word comes from the top bits of h,
// a short literal's bit from the low 5 bits of h, a long literal's bit from the low 5 bits of the byte before
// its K-tail (both are what a funnel shift by that register looks at: no masking in the kernel)
HSB_HD u32 tail_hash(u32 ktail) { return ktail * 0x9E3779B1u; }
HSB_HD u32 tail_word(u32 h, u32 BL) { return h >> (37 - BL); }            // top BL - 5 bits
HSB_HD u32 tail_short_mask(u32 h) { return 1u << (h & 31); }
HSB_HD u32 tail_long_mask(u32 extra) { return 1u << (extra & 31); }

// ---- first tier of k_turbo2 (turbo2.cuh): "gram table", 2^G1 rows of ONE 32-bit set ------------------------------
// A K-gram (K text bytes ending at some position) is looked up as  row = gram_row(first K-1 bytes), bit = low 5 bits of
// a fifth quantity.  The row hash is a multiplication of the 32-bit text window that starts at the gram's first byte by
// GRAM_MULT << 8*(5-K): the shift pushes the window bytes that are NOT among the K-1 hashed ones out of the product, so
// the kernel needs no masking (one IMAD + one SHF per lookup), and the low 8*(K-1) bits of x * odd are a bijection of
// the hashed bytes, so rows are used evenly.
//   Two neighbouring end positions e-1, e share the K-1 bytes S = text[e-K+1 .. e-1]; the row of S holds the union of
//   A = {low5(x) : x S is some literal's K-tail}  and  B = {low5(y) : S y is}; ONE 4-byte load filters both positions
//   (position e-1 tests low5(text[e-K]), position e tests low5(text[e])).
constexpr u32 GRAM_MULT = 0x9E3779B1u;
HSB_HD u32 gram_mul(u32 K) { return K >= 5 || K <= 1 ? (K <= 1 ? 0u : GRAM_MULT) : GRAM_MULT << (8 * (5 - K)); }
HSB_HD u32 gram_row(u32 window, u32 mul, u32 G1) { return (window * mul) >> (32 - G1); }  // G1 >= 10

struct Tables {
    u32 K = 0, H1 = 0, H2 = 0;
    std::vector<u64> pair;   // low 32 bits = A, high 32 bits = B
    // k_turbo2
    u32 G1 = 0, gmul = 0;
    u32 wide = 0;            // 1: this set is better served by the pair table's 8-byte rows (k_turbo)
    std::vector<u32> gram;
    double est_tests_per_byte = 0;  // second-tier tests per text byte expected on uniformly random text
    std::vector<Slot> slots;
    u32 L = 0, BL = 0;       // long-tail length K + 1 (0: every literal is K bytes long) and log2(tail table bits)
    std::vector<u32> bloom;
    u64 variants = 0;
    bool all_simple = true;  // every literal's program pre-decoded to a plain report (no interpreter needed)
};

// Decode the shapes  [CHECK_MED_LIT*] (REPORT|FINAL_REPORT) [END]  with offset_adjust 0.
inline void predecode(const DbView &d, u32 prog, Slot &s, u32 size) {
    u32 kind = SLOT_PROGRAM, target = prog, lit_off = 0, lit_len = 0;
    size_t pc = prog;
    u8 code = rd8(d.bc, pc);
    bool has_lit = false, nocase = false, ok = true;
    if (code == op::CHECK_MED_LIT || code == op::CHECK_MED_LIT_NOCASE) {
        has_lit = true;
        nocase = code == op::CHECK_MED_LIT_NOCASE;
        lit_off = rd32(d.bc, pc + op::check_lit::lit_offset);
        lit_len = rd32(d.bc, pc + op::check_lit::lit_length);
        const u32 fail = rd32(d.bc, pc + op::check_lit::fail_jump);
        // the failure target must be END (nothing else would run)
        ok = lit_len < 65536 && in_range(d, pc + fail, 8) && rd8(d.bc, pc + fail) == op::END;
        pc += op::check_lit::SIZE;
        code = rd8(d.bc, pc);
    }
    if (ok && (code == op::REPORT || code == op::FINAL_REPORT) &&
        rd32(d.bc, pc + op::report::offset_adjust) == 0 &&
        (code == op::FINAL_REPORT || rd8(d.bc, pc + op::report::SIZE) == op::END)) {
        kind = has_lit ? (nocase ? SLOT_LIT_NOCASE_REPORT : SLOT_LIT_REPORT) : SLOT_REPORT;
        target = rd32(d.bc, pc + op::report::onmatch);
    } else {
        lit_off = lit_len = 0;
    }
    s.target = target;
    s.lit_off = lit_off;
    s.meta = size | (kind << 8) | (lit_len << 16);
}

// every n-byte text tail the LitInfo admits (its last n bytes under msk), first tail byte lowest
inline bool tail_variants(const LitRec &l, u32 n, size_t limit, std::vector<u64> &keys) {
    keys.assign(1, 0);
    for (u32 j = 0; j < n; j++) {
        const u32 sh = 8 * (8 - n + j);
        const u8 m = (u8)(l.msk >> sh), val = (u8)(l.v >> sh);
        std::vector<u64> next;
        for (u32 c = 0; c < 256; c++)
            if ((c & m) == val)
                for (u64 k : keys) next.push_back(k | ((u64)c << (8 * j)));
        if (next.empty() || next.size() > limit) return false;
        keys.swap(next);
    }
    return true;
}

inline bool build(const DbView &d, Tables &t) {
    std::vector<LitRec> lits = d.lits;
    if (d.kind == ENG_NOODLE) {  // the single literal, brought to LitInfo form
        LitRec lr{};
        const u32 sh = 8 * (8 - d.nood_len);
        lr.v = d.nood_cmp << sh;
        lr.msk = d.nood_msk << sh;
        lr.groups = ~0ull;
        lr.program = d.nood_id;
        lr.size = (u8)d.nood_len;
        lits.push_back(lr);
    }
    u32 K = 4;
    std::vector<const LitRec *> live;
    for (const LitRec &l : lits) {
        if (!(l.groups & (d.kind == ENG_NOODLE ? ~0ull : d.groups))) continue;
        if (l.size < 1 || l.size > 8) return false;
        K = std::min<u32>(K, l.size);
        live.push_back(&l);
    }
    if (live.empty()) return false;
    t.K = K;
    // enumerate tail variants
    struct Var { u32 key; const LitRec *lit; };
    std::vector<Var> vars;
    for (const LitRec *l : live) {
        std::vector<u32> keys{0};
        for (u32 j = 0; j < K; j++) {
            const u32 sh = 8 * (8 - K + j);
            const u8 m = (u8)(l->msk >> sh), val = (u8)(l->v >> sh);
            std::vector<u32> next;
            for (u32 c = 0; c < 256; c++)
                if ((c & m) == val)
                    for (u32 k : keys) next.push_back(k | (c << (8 * j)));
            if (next.empty() || next.size() > 4096) return false;
            keys.swap(next);
        }
        for (u32 k : keys) vars.push_back({k, l});
        if (vars.size() > (1u << 22)) return false;
    }
    t.variants = vars.size();
    t.H1 = (K == 3 || vars.size() >= 512) ? 14 : (vars.size() >= 64 ? 12 : 10);
    t.pair.assign((size_t)1 << t.H1, 0);
    for (const Var &v : vars) {
        u32 b[4] = {0, 0, 0, 0};
        for (u32 j = 0; j < K; j++) b[j] = (v.key >> (8 * j)) & 0xff;
        // A: extra byte first (b[0]), shared = b[1..K-1];  B: shared = b[0..K-2], extra byte last (b[K-1])
        const u32 sa = pack_shared(K, b[1], b[2], b[3]), sb = pack_shared(K, b[0], b[1], b[2]);
        t.pair[pair_row(K, sa, t.H1)] |= 1ull << (31 - (b[0] & 31));  // sets are stored bit-reversed (turbo.cuh)
        t.pair[pair_row(K, sb, t.H1)] |= 1ull << (63 - (b[K - 1] & 31));
    }
    // tail table
    {
        u32 longest = 0;
        for (const LitRec *l : live) longest = std::max<u32>(longest, l->size);
        t.L = longest > K ? K + 1 : 0;
        struct Entry { u32 ktail; u32 extra; bool is_long; };
        std::vector<Entry> entries;
        std::vector<u64> keys;
        for (const LitRec *l : live) {
            if (t.L && l->size >= t.L && tail_variants(*l, t.L, 256, keys)) {
                for (u64 k : keys) entries.push_back({(u32)(k >> 8), (u32)(k & 0xff), true});
            } else {
                if (!tail_variants(*l, K, 4096, keys)) return false;
                for (u64 k : keys) entries.push_back({(u32)k, 0, false});
            }
        }
        // 2^18 bits (32 KiB) at most: with the 128 KiB pair table and the per-warp text buffers and queues it
        // has to fit the 227 KiB of shared memory an SM offers
        u32 BL = 13;
        while (BL < 18 && ((size_t)1 << BL) < entries.size() * 32) BL++;
        t.BL = BL;
        t.bloom.assign((size_t)1 << (BL - 5), 0);
        for (const Entry &e : entries) {
            const u32 h = tail_hash(e.ktail);
            t.bloom[tail_word(h, BL)] |= e.is_long ? tail_long_mask(e.extra) : tail_short_mask(h);
        }
    }
    // gram table of k_turbo2
    {
        const u32 mul = gram_mul(K);
        t.gmul = mul;
        // rows: a few hundred cells per filed gram, at most 2^15 rows (128 KiB of shared memory)
        u32 G1 = 10;
        while (G1 < 15 && ((size_t)32 << G1) < vars.size() * 400) G1++;
        t.G1 = G1;
        const double cells = 32.0 * (double)((size_t)1 << G1);
        auto window_of = [&](const u32 *b, u32 n) { u32 w = 0; for (u32 j = 0; j < n && j < 4; j++) w |= b[j] << (8 * j); return w; };
        // stride 1: the K-tail variants, A-type and B-type
        std::vector<u32> g1((size_t)1 << G1, 0);
        for (const Var &v : vars) {
            u32 b[4] = {0, 0, 0, 0};
            for (u32 j = 0; j < K; j++) b[j] = (v.key >> (8 * j)) & 0xff;
            g1[gram_row(window_of(b + 1, K - 1), mul, G1)] |= 0x80000000u >> (b[0] & 31);      // A: shared = b[1..K-1]
            g1[gram_row(window_of(b, K - 1), mul, G1)] |= 0x80000000u >> (b[K - 1] & 31);      // B: shared = b[0..K-2]
        }
        // Share of positions the union form lets through, estimated for text over the literals' own alphabet: the text's
        // K-1 shared bytes reach at most alphabet^(K-1) rows however large the table is.
        double tests1;
        {
            bool seen[256] = {false};
            for (const Var &v : vars) for (u32 j = 0; j < K; j++) seen[(v.key >> (8 * j)) & 0xff] = true;
            double alpha = 0, reach = 1, bits = 0, populated = 0;
            for (bool b : seen) alpha += b ? 1 : 0;
            for (u32 j = 1; j < K; j++) reach *= alpha;
            for (u32 w : g1) { bits += __builtin_popcount(w); populated += w ? 1 : 0; }
            reach = std::min(reach, cells / 32.0);
            tests1 = bits / (32.0 * std::max(reach, populated));
        }
        // When the union sets would pass more than a few percent of all positions (many literals behind a SHORT shared
        // part: K = 3 hashes two bytes, so printable text reaches < 10^4 rows however large the table), the 8-byte rows
        // of the pair table (separate A and B sets, k_turbo) halve the candidates the later tiers have to look at and
        // win despite the costlier lookup (measured on C2: DESIGN.md section 4).
        const char *force = getenv("HS_GPU_TURBO_WIDE");
        const int forced = force ? atoi(force) : -1;
        t.wide = (K >= 2 && K <= 3 && (forced == 1 || (forced != 0 && tests1 > 0.04))) ? 1 : 0;
        t.gram.swap(g1);
        t.est_tests_per_byte = tests1;
    }
    u32 H2 = 4;
    while (((size_t)1 << H2) < vars.size() * 4) H2++;
    if (H2 > 26) return false;
    t.H2 = H2;
    t.slots.assign((size_t)1 << H2, Slot{0, 0, 0, 0, 0, 0});
    const u32 mask = (1u << H2) - 1;
    // slots of one key are filed in ascending report id, so a probe walk meets them in (id) order
    std::vector<Slot> filed;
    filed.reserve(vars.size());
    for (const Var &v : vars) {
        Slot s{v.lit->v, v.lit->msk, v.key, 0, 0, 0};
        predecode(d, v.lit->program, s, v.lit->size);
        if (((s.meta >> 8) & 0xf) == SLOT_PROGRAM) t.all_simple = false;
        filed.push_back(s);
    }
    std::stable_sort(filed.begin(), filed.end(), [](const Slot &a, const Slot &b) {
        return a.key != b.key ? a.key < b.key : a.target < b.target;
    });
    for (const Slot &s : filed) {
        u32 h = slot_home(s.key, H2);
        while ((t.slots[h].meta >> 8) & 0xf) h = (h + 1) & mask;
        t.slots[h] = s;
    }
    return true;
}

}  // namespace turbo
}  // namespace hsb
