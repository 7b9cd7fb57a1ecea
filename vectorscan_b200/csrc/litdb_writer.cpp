// litdb_writer.cpp -- host-side writer of PURE-LITERAL block-mode databases in the Vectorscan 5.4.11
// bytecode format (libhs_litdb.so).
//
// Why this exists: the north star keeps the reference's host compiler (hs_compile_multi) unchanged,
// but that compiler needs Boost + Ragel and cannot be built in this image.
// For literal sets this file stands in for hs_compile_lit_multi (/root/reference/src/hs.cpp:437-445):
// it emits the same relocatable blob -- hs_database header, RoseEngine (runtimeImpl = PURE_LITERAL),
// one Rose program per literal fragment, an HWLM table holding a noodle / Teddy / FDR engine with
// confirm and flood structures -- so that the UNMODIFIED reference runtime accepts it (dbIsValid)
// and scans it with hs_scan.  The GPU runtime reads the very same bytes.  Correctness criterion
// (tests/test_litdb_vs_reference.py): reference hs_scan on the emitted blob == brute-force matches.
//
// Format/algorithm sources followed (restated, nothing copied):
//   table choice      src/hwlm/hwlm_build.cpp:106-118,201-256, src/fdr/fdr_engine_description.cpp:62-199,
//                     src/fdr/teddy_engine_description.cpp:90-190
//   FDR table         src/fdr/fdr_compile.cpp:129-211 (layout, start state), :385-511 (bucket DP), :526-631 (table)
//   confirm           src/fdr/fdr_confirm_compile.cpp:72-338
//   flood             src/fdr/flood_compile.cpp:92-230
//   Teddy             src/fdr/teddy_compile.cpp:97-314 (packing), :318-557 (masks, reinforcement), :559-620 (layout)
//   noodle            src/hwlm/noodle_build.cpp:66-140
//   Rose programs     src/rose/rose_build_program.cpp:534-640,788-823 ; engine header src/rose/rose_build_bytecode.cpp:381-464,3608-3890
//   database header   src/database.h:103-114, src/database.c:88-106,178-197
#include "hsb_layout.h"
#include "hs_litdb.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <map>
#include <set>
#include <string>
#include <vector>

namespace {
using namespace hsb;

// ---- CRC32C (Castagnoli, reflected 0x82f63b78, zero init, no final xor: src/crc32.c:589-596) ----
u32 crc32c(const u8 *p, size_t n) {
    static u32 table[256];
    static bool init = false;
    if (!init) {
        for (u32 i = 0; i < 256; i++) {
            u32 c = i;
            for (int k = 0; k < 8; k++) c = (c & 1) ? (c >> 1) ^ 0x82f63b78u : c >> 1;
            table[i] = c;
        }
        init = true;
    }
    u32 crc = 0;
    for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xff] ^ (crc >> 8);
    return crc;
}

inline bool is_alpha(u8 c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z'); }
inline u8 to_upper(u8 c) { return (c >= 'a' && c <= 'z') ? (u8)(c - 0x20) : c; }
inline u8 to_lower(u8 c) { return (c >= 'A' && c <= 'Z') ? (u8)(c + 0x20) : c; }

struct Buf {
    std::vector<u8> b;
    size_t size() const { return b.size(); }
    void align(size_t a) { b.resize(round_up(b.size(), a), 0); }
    size_t grow(size_t n) { size_t o = b.size(); b.resize(o + n, 0); return o; }
    u8 *at(size_t off) { return b.data() + off; }
};

// One literal as handed to the literal matcher (hwlmLiteral, src/hwlm/hwlm_literal.h): at most 8
// bytes (the suffix of the full literal), one nocase flag, id = Rose program offset.
struct HwlmLit {
    std::string s;
    bool nocase;
    u32 id;
    u64 groups;
};

struct FullLit {
    std::string s;
    bool nocase;
    u32 report;
    bool single;  // HS_FLAG_SINGLEMATCH
    u32 ekey = INVALID_EKEY;
    u32 dkey = INVALID_DKEY;
};

// ------------------------------------------------------------------------------------------------
// confirm structures: u32 confBase[nBuckets] (cacheline rounded) then one FDRConfirm per bucket
// ------------------------------------------------------------------------------------------------
void lit_value_mask(const HwlmLit &l, u64 &v, u64 &msk) {
    // little-endian 8-byte window ending at the literal's last byte; last byte in the top byte
    v = 0;
    msk = 0;
    for (size_t j = 0; j < 8 && j < l.s.size(); j++) {
        u8 c = (u8)l.s[l.s.size() - 1 - j];
        unsigned sh = (unsigned)(7 - j) * 8;
        if (l.nocase && is_alpha(c)) {
            msk |= (u64)0xdf << sh;
            v |= (u64)(c & 0xdf) << sh;
        } else {
            msk |= (u64)0xff << sh;
            v |= (u64)c << sh;
        }
    }
}

u32 floor_log2(u32 x) { u32 r = 0; while (x >>= 1) r++; return r; }

std::vector<u8> build_bucket_confirm(const std::vector<HwlmLit> &lits) {
    const size_t n = lits.size();
    std::vector<u64> v(n), m(n);
    u64 andmsk = ~0ull, groups = 0;
    for (size_t i = 0; i < n; i++) {
        lit_value_mask(lits[i], v[i], m[i]);
        andmsk &= m[i];
        groups |= lits[i].groups;
    }
    const u32 nBits = floor_log2((u32)n) + 4;
    std::map<u32, std::vector<size_t>> by_hash;
    for (size_t i = 0; i < n; i++) {
        u32 h = (u32)(((v[i] & andmsk) * conf::MULT) >> (64 - nBits));
        by_hash[h].push_back(i);
    }
    Buf out;
    out.grow(conf::SIZE);
    wr64(out.at(0), conf::andmsk, andmsk);
    wr64(out.at(0), conf::mult, conf::MULT);
    wr32(out.at(0), conf::nBits, nBits);
    wr64(out.at(0), conf::groups, groups);
    size_t index_off = out.grow((size_t)4 << nBits);
    out.align(8);
    for (auto &kv : by_hash) {
        wr32(out.at(index_off), (size_t)kv.first * 4, (u32)out.size());
        for (size_t k = 0; k < kv.second.size(); k++) {
            size_t i = kv.second[k];
            size_t o = out.grow(lit::SIZE);
            wr64(out.at(o), lit::v, v[i]);
            wr64(out.at(o), lit::msk, m[i]);
            wr64(out.at(o), lit::groups, lits[i].groups);
            wr32(out.at(o), lit::id, lits[i].id);
            wr8(out.at(o), lit::size, (u8)lits[i].s.size());
            wr8(out.at(o), lit::flags, 0);
            wr8(out.at(o), lit::next, k + 1 < kv.second.size() ? 1 : 0);
        }
    }
    out.align(8);
    return out.b;
}

std::vector<u8> build_confirm(const std::vector<HwlmLit> &lits,
                              const std::vector<std::vector<u32>> &buckets) {
    Buf out;
    out.grow(round_up(buckets.size() * 4, CACHELINE));
    for (size_t b = 0; b < buckets.size(); b++) {
        if (buckets[b].empty()) continue;
        std::vector<HwlmLit> bl;
        for (u32 i : buckets[b]) bl.push_back(lits[i]);
        std::vector<u8> fc = build_bucket_confirm(bl);
        wr32(out.at(0), b * 4, (u32)out.size());
        size_t o = out.grow(fc.size());
        std::memcpy(out.at(o), fc.data(), fc.size());
    }
    return out.b;
}

// ------------------------------------------------------------------------------------------------
// flood control: u32 index[256] then the distinct FDRFlood records
// ------------------------------------------------------------------------------------------------
struct FloodRec {
    u64 allGroups = 0;
    u32 suffix = 0;
    u16 idCount = 0;
    u32 ids[flood::MAX_IDS] = {0};
    u64 groups[flood::MAX_IDS] = {0};
    std::vector<u8> bytes() const {
        std::vector<u8> r(flood::SIZE, 0);
        wr64(r.data(), flood::allGroups, allGroups);
        wr32(r.data(), flood::suffix, suffix);
        wr16(r.data(), flood::idCount, idCount);
        for (u32 i = 0; i < flood::MAX_IDS; i++) {
            wr32(r.data(), flood::ids + 4 * i, ids[i]);
            wr64(r.data(), flood::groups + 8 * i, groups[i]);
        }
        return r;
    }
};

std::vector<u8> build_flood(const std::vector<HwlmLit> &lits, u32 default_suffix, bool allow_flood) {
    std::vector<FloodRec> per_char(256);
    for (auto &f : per_char) f.suffix = default_suffix;
    auto bump = [&](u8 c, u32 run) { per_char[c].suffix = std::max(per_char[c].suffix, run + 1); };
    auto add = [&](u8 c, const HwlmLit &l, u32 run) {
        FloodRec &f = per_char[c];
        f.suffix = std::max(f.suffix, run + 1);
        if (f.idCount < flood::MAX_IDS) {
            f.ids[f.idCount] = l.id;
            f.groups[f.idCount] = l.groups;
            f.allGroups |= l.groups;
            f.idCount++;
        }
    };
    for (const HwlmLit &l : lits) {
        const u32 len = (u32)l.s.size();
        const u8 last = (u8)l.s[len - 1];
        const bool fold = l.nocase && is_alpha(last);
        // length of the trailing run of `last` (case-folded when the literal is caseless)
        u32 run = len;
        for (u32 i = 0; i < len; i++) {
            u8 c = (u8)l.s[len - 1 - i];
            bool same = l.nocase ? to_lower(c) == to_lower(last) : c == last;
            if (!same) { run = i; break; }
        }
        const bool whole = run == len;  // the literal is a pure run: a flood of `last` fires it per byte
        u8 c0 = fold ? to_upper(last) : last;
        if (whole) add(c0, l, run); else bump(c0, run);
        if (fold) { if (whole) add(to_lower(last), l, run); else bump(to_lower(last), run); }
    }
    if (!allow_flood) for (auto &f : per_char) f.idCount = flood::MAX_IDS;
    std::map<std::vector<u8>, u32> distinct;
    std::vector<std::vector<u8>> recs;
    std::vector<u32> index(256);
    // records are laid out in memcmp order of their bytes, as a std::map keyed on the struct would
    std::set<std::vector<u8>> ordered;
    for (int c = 0; c < 256; c++) ordered.insert(per_char[c].bytes());
    for (auto &r : ordered) { distinct[r] = (u32)recs.size(); recs.push_back(r); }
    for (int c = 0; c < 256; c++) index[c] = distinct[per_char[c].bytes()];
    Buf out;
    out.grow(256 * 4);
    for (int c = 0; c < 256; c++) wr32(out.at(0), 4 * c, index[c]);
    for (auto &r : recs) { size_t o = out.grow(r.size()); std::memcpy(out.at(o), r.data(), r.size()); }
    out.align(16);
    return out.b;
}

// ------------------------------------------------------------------------------------------------
// FDR
// ------------------------------------------------------------------------------------------------
struct FdrShape { u32 domain; u32 stride; };

FdrShape choose_fdr_shape(const std::vector<HwlmLit> &lits) {
    size_t min_len = ~(size_t)0, min_count = 0;
    for (auto &l : lits) {
        if (l.s.size() < min_len) { min_len = l.s.size(); min_count = 0; }
        if (l.s.size() == min_len) min_count++;
    }
    const size_t n = lits.size();
    u32 want = 1;
    if (min_len > 1) {
        if (n < 250) want = (u32)min_len;
        else if (n < 800) want = (u32)min_len - 1;
        else if (n < 5000) want = std::min<u32>((u32)min_len - 1, 2);
    }
    if (min_len == 4 && want == 4 && min_count > 2) want = 2;
    FdrShape best{0, 0};
    int best_score = -1;
    for (u32 domain = 9; domain <= 15; domain++) {
        for (u32 stride = 1; stride <= 4; stride *= 2) {
            if (domain > 13 && stride > 1) continue;
            if (min_len < stride) continue;
            int score = 100 - std::abs((int)want - (int)stride);
            if (stride <= want) score += (int)stride;
            int ideal;
            if (n < 8) ideal = stride == 1 ? 8 : 10;
            else if (n < 20) ideal = 10;
            else if (n < 100) ideal = 11;
            else if (n < 1000) ideal = 12;
            else if (n < 10000) ideal = 13;
            else ideal = 15;
            if (stride > 1) ideal++;
            score -= std::abs(ideal - (int)domain);
            if (score > best_score) { best_score = score; best = {domain, stride}; }
        }
    }
    return best;
}

// Sort by length then by reversed text, cut into same-length chunks, and choose the 8 bucket
// boundaries minimising sum(count^1.05 / len^3) with a DP over chunk boundaries.
std::vector<std::vector<u32>> assign_fdr_buckets(std::vector<HwlmLit> &lits) {
    std::stable_sort(lits.begin(), lits.end(), [](const HwlmLit &a, const HwlmLit &b) {
        if (a.s.size() != b.s.size()) return a.s.size() < b.s.size();
        for (size_t i = 0; i < a.s.size(); i++) {
            u8 ca = (u8)a.s[a.s.size() - 1 - i], cb = (u8)b.s[b.s.size() - 1 - i];
            if (ca != cb) return (signed char)ca < (signed char)cb;
        }
        return a.nocase > b.nocase;
    });
    const u32 n = (u32)lits.size();
    std::set<size_t> lens;
    for (auto &l : lits) lens.insert(l.s.size());
    const u32 CHUNK_MAX = 512;
    const u32 per_chunk = n / (CHUNK_MAX - (u32)std::min<size_t>(16, lens.size())) + 1;
    struct Chunk { u32 first, count, len; };
    std::vector<Chunk> chunks;
    for (u32 i = 0; i < n; i++) {
        bool same_text = i && lits[i].s.size() == lits[i - 1].s.size() && [&] {
            for (size_t k = 0; k < lits[i].s.size(); k++)
                if (to_upper((u8)lits[i].s[k]) != to_upper((u8)lits[i - 1].s[k])) return false;
            return true;
        }();
        bool cut = chunks.empty() || (!same_text && chunks.size() < CHUNK_MAX - 1 &&
                   (lits[i].s.size() != chunks.back().len || i - chunks.back().first >= per_chunk));
        if (cut) chunks.push_back({i, 0, (u32)lits[i].s.size()});
        chunks.back().count++;
    }
    const u32 C = (u32)chunks.size(), B = fdr::NUM_BUCKETS;
    auto cost = [&](u32 len, u32 count) { return std::pow((double)count, 1.05) / std::pow((double)len, 3.0); };
    std::vector<u32> prefix(C + 1, 0);
    for (u32 j = 0; j < C; j++) prefix[j + 1] = prefix[j] + chunks[j].count;
    const double INF = 1e300;
    // best[i][j]: cheapest way to place chunks j..C-1 into at most i+1 buckets; a bucket is scored
    // by the length of its shortest (first) chunk
    std::vector<std::vector<double>> best(B, std::vector<double>(C + 1, INF));
    std::vector<std::vector<u32>> cutpt(B, std::vector<u32>(C + 1, C));
    for (u32 i = 0; i < B; i++) best[i][C] = 0;
    for (u32 j = 0; j < C; j++) best[0][j] = cost(chunks[j].len, prefix[C] - prefix[j]);
    for (u32 i = 1; i < B; i++) {
        for (u32 j = 0; j < C; j++) {
            best[i][j] = best[i - 1][j];
            cutpt[i][j] = cutpt[i - 1][j];
            for (u32 k = j + 1; k <= C; k++) {
                double s = cost(chunks[j].len, prefix[k] - prefix[j]) + best[i - 1][k];
                if (s < best[i][j]) { best[i][j] = s; cutpt[i][j] = k; }
            }
        }
    }
    std::vector<std::vector<u32>> groups;  // shortest literals first
    u32 j = 0;
    for (int i = (int)B - 1; j < C && i >= 0; i--) {
        u32 k = i == 0 ? C : cutpt[i][j];
        std::vector<u32> ids;
        for (u32 q = prefix[k]; q-- > prefix[j];) ids.push_back(q);  // longest first inside a bucket
        groups.push_back(ids);
        j = k;
    }
    // bucket 0 holds the longest literals
    std::vector<std::vector<u32>> buckets(B);
    for (size_t g = 0; g < groups.size(); g++) buckets[groups.size() - 1 - g] = groups[g];
    return buckets;
}

std::vector<u8> build_fdr(std::vector<HwlmLit> lits, FdrShape shape, bool allow_flood) {
    std::vector<std::vector<u32>> buckets = assign_fdr_buckets(lits);
    const u32 domain = shape.domain, rows = 1u << domain, dmask = rows - 1;
    std::vector<u64> tab(rows, ~0ull);
    u64 always_clear = 0;  // (bucket, pos) pairs whose probe is unconstrained for every row
    for (u32 b = 0; b < fdr::NUM_BUCKETS; b++) {
        for (u32 pos = 0; pos < 8; pos++) {
            const u64 bit = 1ull << (pos * 8 + b);
            bool unconstrained = false;
            // (dontcare mask -> set of values) for the two text bytes a probe at this suffix
            // position sees: low byte = literal byte `pos` from the end, high byte = the one after it
            std::map<u32, std::set<u32>> pats;
            for (u32 li : buckets[b]) {
                const HwlmLit &l = lits[li];
                const int sz = (int)l.s.size();
                u32 val = 0, dc = 0;
                for (int k = 0; k < 2; k++) {
                    int p = (int)pos - k;
                    if (p < 0 || p >= sz) { dc |= 0xffu << (8 * k); continue; }
                    u8 c = (u8)l.s[sz - 1 - p];
                    u8 vb = c, db = 0;
                    int remain = (int)domain - 8 * k;
                    if (remain < 8) { u8 keep = (u8)((1u << remain) - 1); vb &= keep; db |= (u8)~keep; }
                    if (l.nocase && is_alpha(c)) { vb &= 0xdf; db |= 0x20; }
                    val |= (u32)vb << (8 * k);
                    dc |= (u32)db << (8 * k);
                }
                val &= dmask;
                dc &= dmask;
                if (dc == dmask) { unconstrained = true; break; }
                pats[dc].insert(val);
            }
            if (unconstrained) { always_clear |= bit; continue; }
            for (auto &kv : pats) {
                const u32 dc = kv.first;
                // enumerate every assignment of the don't-care bits
                u32 sub = 0;
                do {
                    for (u32 v : kv.second) tab[(v & ~dc) | sub] &= ~bit;
                    sub = (sub - dc) & dc;
                } while (sub);
            }
        }
    }
    for (auto &r : tab) r &= ~always_clear;

    std::vector<u8> confirm = build_confirm(lits, buckets);
    std::vector<u8> floodt = build_flood(lits, /* schemeWidth 64 / 8 buckets + 1 */ 9, allow_flood);
    const size_t tab_bytes = (size_t)rows * 8;
    const size_t conf_off = fdr::TABLE + round_up(tab_bytes, CACHELINE);
    const size_t flood_off = conf_off + round_up(confirm.size(), CACHELINE);
    const size_t total = flood_off + floodt.size();
    std::vector<u8> out(total, 0);
    u32 max_len = 0;
    for (auto &l : lits) max_len = std::max<u32>(max_len, (u32)l.s.size());
    wr32(out.data(), fdr::engineID, 0);
    wr32(out.data(), fdr::size, (u32)total);
    wr32(out.data(), fdr::maxStringLen, max_len);
    wr32(out.data(), fdr::numStrings, (u32)lits.size());
    wr32(out.data(), fdr::confOffset, (u32)conf_off);
    wr32(out.data(), fdr::floodOffset, (u32)flood_off);
    wr8(out.data(), fdr::stride, (u8)shape.stride);
    wr8(out.data(), fdr::domain, (u8)domain);
    wr16(out.data(), fdr::domainMask, (u16)dmask);
    wr32(out.data(), fdr::tabSize, (u32)tab_bytes);
    // start state: for each bucket, end positions < (shortest literal - 1) cannot be candidates
    for (u32 b = 0; b < fdr::NUM_BUCKETS; b++) {
        u32 shortest = ~0u;
        for (u32 li : buckets[b]) shortest = std::min<u32>(shortest, (u32)lits[li].s.size());
        for (u32 pos = 0; pos < 8; pos++)
            if (shortest != ~0u ? pos + 1 < shortest : true) out[fdr::start + pos] |= (u8)(1u << b);
    }
    std::memcpy(out.data() + fdr::TABLE, tab.data(), tab_bytes);
    std::memcpy(out.data() + conf_off, confirm.data(), confirm.size());
    std::memcpy(out.data() + flood_off, floodt.data(), floodt.size());
    return out;
}

// ------------------------------------------------------------------------------------------------
// Teddy (8 buckets, 1..4 masks)
// ------------------------------------------------------------------------------------------------
struct NibbleSet {
    std::vector<u16> sets;  // per mask: lo-nibble set, hi-nibble set
    std::vector<u32> lits;
    u64 weight() const {
        u64 p = 1;
        for (u16 s : sets) p *= (u64)__builtin_popcount(s);
        return p * (2 + lits.size());
    }
    bool flood_prone(u32 masks) const {
        u16 lo = 0xffff, hi = 0xffff;
        for (u32 i = 0; i < masks; i++) { lo &= sets[2 * i]; hi &= sets[2 * i + 1]; }
        return lo && hi;
    }
};

NibbleSet nibbles_of(u32 idx, const HwlmLit &l, u32 masks) {
    NibbleSet ns;
    ns.sets.assign(2 * masks, 0xffff);
    for (u32 i = 0; i < masks && i < l.s.size(); i++) {
        u8 c = (u8)l.s[l.s.size() - 1 - i];
        ns.sets[2 * i] = (u16)(1u << (c & 0xf));
        u8 hi = c >> 4;
        ns.sets[2 * i + 1] = (l.nocase && is_alpha(c)) ? (u16)((1u << (hi & 0xd)) | (1u << (hi | 0x2)))
                                                        : (u16)(1u << hi);
    }
    ns.lits.push_back(idx);
    return ns;
}

bool pack_teddy(const std::vector<HwlmLit> &lits, u32 masks, u32 nbuckets,
                std::vector<std::vector<u32>> &buckets) {
    std::vector<NibbleSet> sets;
    for (u32 i = 0; i < lits.size(); i++) sets.push_back(nibbles_of(i, lits[i], masks));
    auto merged = [](const NibbleSet &a, const NibbleSet &b) {
        NibbleSet m = a;
        for (size_t i = 0; i < m.sets.size(); i++) m.sets[i] |= b.sets[i];
        m.lits.insert(m.lits.end(), b.lits.begin(), b.lits.end());
        std::sort(m.lits.begin(), m.lits.end());
        return m;
    };
    for (;;) {
        int bi = -1, bj = -1;
        u64 best = ~0ull;
        bool free_win = false;
        for (size_t i = 0; i < sets.size() && !free_win; i++) {
            for (size_t j = i + 1; j < sets.size(); j++) {
                if (sets.size() <= nbuckets && sets[i].sets != sets[j].sets) continue;
                NibbleSet m = merged(sets[i], sets[j]);
                u64 after = m.weight(), before = sets[i].weight() + sets[j].weight();
                if (after < before) { bi = (int)i; bj = (int)j; free_win = true; break; }
                if (m.flood_prone(masks) && !(sets[i].flood_prone(masks) && sets[j].flood_prone(masks))) continue;
                if (after - before < best) { best = after - before; bi = (int)i; bj = (int)j; }
            }
        }
        if (bi < 0) break;
        NibbleSet m = merged(sets[bi], sets[bj]);
        sets.erase(sets.begin() + bj);
        sets.erase(sets.begin() + bi);
        sets.push_back(m);
    }
    if (sets.size() > nbuckets) return false;
    buckets.assign(nbuckets, {});
    for (size_t b = 0; b < sets.size(); b++) buckets[b] = sets[b].lits;
    return true;
}

u32 choose_teddy_masks(const std::vector<HwlmLit> &lits) {
    // 0 = Teddy not applicable.  Restates the 8-bucket rows of chooseTeddyEngine's scoring.
    if (lits.size() > 48) return 0;
    size_t max_len = 0, flood_tail = 0;
    for (auto &l : lits) {
        max_len = std::max(max_len, l.s.size());
        size_t j = 1;
        while (j < l.s.size() && l.s[l.s.size() - 1 - j] == l.s.back()) j++;
        flood_tail = std::max(flood_tail, j);
    }
    u32 best = 0;
    int best_score = -1;
    for (u32 packed = 0; packed < 2; packed++) {
        for (u32 masks = 1; masks <= 4; masks++) {
            if (!packed && lits.size() > 8) continue;
            if (masks > max_len) continue;
            if (lits.size() > 40) {
                size_t small = 0;
                for (auto &l : lits) small += l.s.size() < masks;
                if (small * 5 > lits.size()) continue;
            }
            int score = packed ? 0 : 100;
            score += lits.size() > 32 ? (int)masks * 4 : 100;
            if (masks > flood_tail) score += 50;
            score += 6 / (std::abs(3 - (int)masks) + 1);
            score += 2;
            if (score > best_score) { best_score = score; best = masks | (packed << 8); }
        }
    }
    return best;
}

std::vector<u8> build_teddy(const std::vector<HwlmLit> &lits, u32 masks, bool packed, bool allow_flood,
                            std::string &err) {
    std::vector<std::vector<u32>> buckets;
    if (!pack_teddy(lits, masks, 8, buckets)) { err = "literal set does not pack into 8 Teddy buckets"; return {}; }
    const size_t mask_len = (size_t)masks * 32;
    const size_t rtab_len = teddy::REINFORCE_ROWS * 8;
    std::vector<u8> nib(mask_len, 0xff);
    std::vector<u64> reinforce(teddy::REINFORCE_ROWS, 0x00ffffffffffffffull);
    for (u32 b = 0; b < 8; b++) {
        const u8 bm = (u8)(1u << b);
        for (u32 li : buckets[b]) {
            const HwlmLit &l = lits[li];
            const u32 sz = (u32)l.s.size();
            for (u32 j = 0; j < masks; j++) {
                u8 *lo = nib.data() + (size_t)(2 * j) * 16, *hi = nib.data() + (size_t)(2 * j + 1) * 16;
                if (j >= sz) {
                    for (int n = 0; n < 16; n++) { lo[n] &= (u8)~bm; hi[n] &= (u8)~bm; }
                    continue;
                }
                u8 c = (u8)l.s[sz - 1 - j];
                lo[c & 0xf] &= (u8)~bm;
                if (l.nocase && is_alpha(c)) {
                    hi[(c >> 4) & 0xd] &= (u8)~bm;
                    hi[(c >> 4) | 0x2] &= (u8)~bm;
                } else {
                    hi[c >> 4] &= (u8)~bm;
                }
            }
            // reinforcement: byte j-1 of row[c] has bit b clear iff a bucket-b literal can have byte c
            // j positions before its end (row index = the byte preceding the vector being scanned)
            for (u32 j = 1; j < 8; j++) {
                auto clear = [&](u32 row) { reinforce[row] &= ~((u64)bm << (8 * (j - 1))); };
                if (j > sz - 1) { for (u32 r = 0; r < 256; r++) clear(r); continue; }
                u8 c = (u8)l.s[sz - 1 - j];
                if (l.nocase && is_alpha(c)) { clear(c & 0xdf); clear(c | 0x20); } else clear(c);
            }
        }
    }
    reinforce[256] = 0;  // "no reinforcement" row
    std::vector<u8> confirm = build_confirm(lits, buckets);
    std::vector<u8> floodt = build_flood(lits, masks, allow_flood);
    const size_t conf_off = teddy::MASKS + round_up(mask_len, CACHELINE) + round_up(rtab_len, CACHELINE);
    const size_t flood_off = conf_off + round_up(confirm.size(), CACHELINE);
    const size_t total = flood_off + floodt.size();
    std::vector<u8> out(total, 0);
    u32 max_len = 0;
    for (auto &l : lits) max_len = std::max<u32>(max_len, (u32)l.s.size());
    wr32(out.data(), fdr::engineID, teddy::ENGINE_FIRST + 2 * (masks - 1) + (packed ? 1 : 0));
    wr32(out.data(), fdr::size, (u32)total);
    wr32(out.data(), fdr::maxStringLen, max_len);
    wr32(out.data(), fdr::numStrings, (u32)lits.size());
    wr32(out.data(), fdr::confOffset, (u32)conf_off);
    wr32(out.data(), fdr::floodOffset, (u32)flood_off);
    std::memcpy(out.data() + teddy::MASKS, nib.data(), mask_len);
    std::memcpy(out.data() + teddy::MASKS + round_up(mask_len, CACHELINE), reinforce.data(), rtab_len);
    std::memcpy(out.data() + conf_off, confirm.data(), confirm.size());
    std::memcpy(out.data() + flood_off, floodt.data(), floodt.size());
    return out;
}

// ------------------------------------------------------------------------------------------------
// noodle (exactly one literal)
// ------------------------------------------------------------------------------------------------
std::vector<u8> build_noodle(const HwlmLit &l) {
    const size_t len = l.s.size();
    u8 msk[8] = {0}, cmp[8] = {0};
    for (size_t i = 0; i < len; i++) {
        u8 c = (u8)l.s[i];
        u8 m = (l.nocase && is_alpha(c)) ? 0xdf : 0xff;
        msk[i] = m;
        cmp[i] = c & m;
    }
    // key = first adjacent pair that differs (case-folded for caseless alphas), else the last pair
    size_t key = 0;
    for (size_t i = 0; i + 1 < len; i++) {
        u8 c = (u8)l.s[i], d = (u8)l.s[i + 1];
        bool diff = (l.nocase && is_alpha(c)) ? to_upper(c) != to_upper(d) : c != d;
        key = i;
        if (diff) break;
    }
    std::vector<u8> out(nood::SIZE, 0);
    wr32(out.data(), nood::id, l.id);
    u64 m64, c64;
    std::memcpy(&m64, msk, 8);
    std::memcpy(&c64, cmp, 8);
    wr64(out.data(), nood::msk, m64);
    wr64(out.data(), nood::cmp, c64);
    wr8(out.data(), nood::msk_len, (u8)len);
    wr8(out.data(), nood::key_offset, (u8)(len - key));
    wr8(out.data(), nood::nocase, l.nocase ? 1 : 0);
    wr8(out.data(), nood::single, len == 1 ? 1 : 0);
    wr8(out.data(), nood::key0, (u8)l.s[key]);
    wr8(out.data(), nood::key1, len == 1 ? 0 : (u8)l.s[key + 1]);
    return out;
}

// ------------------------------------------------------------------------------------------------
// Rose programs + engine header + database header
// ------------------------------------------------------------------------------------------------
struct Fragment {
    std::string suffix;  // <= 8 bytes handed to the literal matcher
    bool nocase;
    std::vector<u32> fulls;  // indices into the FullLit array
    u32 program = 0;
};

int build(const std::vector<FullLit> &fulls_in, const hsb_lit_options &opt, std::vector<u8> &image,
          std::string &err) {
    std::vector<FullLit> fulls = fulls_in;
    if (fulls.empty()) { err = "no literals"; return HSB_LIT_COMPILER_ERROR; }
    for (auto &f : fulls)
        if (f.s.empty()) { err = "empty literal"; return HSB_LIT_COMPILER_ERROR; }

    // exhaustion keys: one per SINGLEMATCH report id; dedupe keys: report ids reachable from more
    // than one distinct literal (the same (id, to) could otherwise be delivered twice)
    std::map<u32, u32> ekeys, dkeys;
    std::map<u32, std::set<std::pair<std::string, bool>>> lits_per_report;
    for (auto &f : fulls) {
        if (f.single) ekeys.emplace(f.report, 0);
        lits_per_report[f.report].insert({f.s, f.nocase});
    }
    { u32 k = 0; for (auto &e : ekeys) e.second = k++; }
    for (auto &r : lits_per_report) if (r.second.size() > 1) { u32 k = (u32)dkeys.size(); dkeys[r.first] = k; }
    bool all_exhaustible = true;
    for (auto &f : fulls) {
        if (f.single) f.ekey = ekeys[f.report]; else all_exhaustible = false;
        auto d = dkeys.find(f.report);
        if (d != dkeys.end()) f.dkey = d->second;
    }

    // fragments: full literals sharing the same (<=8-byte suffix, caselessness)
    std::map<std::pair<std::string, bool>, u32> frag_index;
    std::vector<Fragment> frags;
    for (u32 i = 0; i < fulls.size(); i++) {
        const FullLit &f = fulls[i];
        std::string suf = f.s.size() > 8 ? f.s.substr(f.s.size() - 8) : f.s;
        bool any_alpha = false;
        for (char c : suf) any_alpha |= is_alpha((u8)c);
        auto key = std::make_pair(suf, f.nocase && any_alpha);
        auto it = frag_index.find(key);
        if (it == frag_index.end()) {
            it = frag_index.emplace(key, (u32)frags.size()).first;
            frags.push_back({suf, key.second, {}, 0});
        }
        // identical (text, case, report) duplicates add nothing
        bool dup = false;
        for (u32 o : frags[it->second].fulls)
            dup |= fulls[o].s == f.s && fulls[o].nocase == f.nocase && fulls[o].report == f.report;
        if (!dup) frags[it->second].fulls.push_back(i);
    }

    Buf rose;
    rose.grow(rose::SIZE);
    rose.align(CACHELINE);

    // inverse dkey table (dkey -> report id)
    u32 inv_dkey_off = 0;
    if (!dkeys.empty()) {
        inv_dkey_off = (u32)rose.grow(4 * dkeys.size());
        for (auto &d : dkeys) wr32(rose.at(inv_dkey_off), 4 * d.second, d.first);
        rose.align(8);
    }

    // literal text for CHECK_MED_LIT
    std::vector<u32> lit_text_off(fulls.size(), 0);
    for (auto &fr : frags)
        for (u32 i : fr.fulls)
            if (fulls[i].s.size() > 8) {
                // caseless literals are stored upper-cased: cmpForward folds only the scanned text
                // (src/util/compare.h:143-175)
                lit_text_off[i] = (u32)rose.grow(fulls[i].s.size());
                for (size_t k = 0; k < fulls[i].s.size(); k++) {
                    u8 c = (u8)fulls[i].s[k];
                    *rose.at(lit_text_off[i] + k) = fulls[i].nocase ? to_upper(c) : c;
                }
            }
    rose.align(8);

    // programs: per full literal a block  [CHECK_MED_LIT*] [CHECK_EXHAUSTED] [DEDUPE] REPORT*  whose
    // fail_jumps land on the next block; END closes the program
    for (auto &fr : frags) {
        rose.align(op::ALIGN);
        fr.program = (u32)rose.size();
        for (size_t bi = 0; bi < fr.fulls.size(); bi++) {
            const FullLit &f = fulls[fr.fulls[bi]];
            size_t block_len = 0;
            if (f.s.size() > 8) block_len += op::check_lit::SIZE;
            if (f.ekey != INVALID_EKEY) {
                block_len += op::check_exhausted::SIZE;
                if (f.dkey != INVALID_DKEY) block_len += op::dedupe::SIZE;
                block_len += op::report_exhaust::SIZE;
            } else if (f.dkey != INVALID_DKEY) {
                block_len += op::dedupe_and_report::SIZE;
            } else {
                block_len += op::report::SIZE;
            }
            const size_t block_start = rose.size(), block_end = block_start + block_len;
            if (f.s.size() > 8) {
                size_t o = rose.grow(op::check_lit::SIZE);
                wr8(rose.at(o), 0, f.nocase ? op::CHECK_MED_LIT_NOCASE : op::CHECK_MED_LIT);
                wr32(rose.at(o), op::check_lit::lit_offset, lit_text_off[fr.fulls[bi]]);
                wr32(rose.at(o), op::check_lit::lit_length, (u32)f.s.size());
                wr32(rose.at(o), op::check_lit::fail_jump, (u32)(block_end - o));
            }
            if (f.ekey != INVALID_EKEY) {
                size_t o = rose.grow(op::check_exhausted::SIZE);
                wr8(rose.at(o), 0, op::CHECK_EXHAUSTED);
                wr32(rose.at(o), op::check_exhausted::ekey, f.ekey);
                wr32(rose.at(o), op::check_exhausted::fail_jump, (u32)(block_end - o));
                if (f.dkey != INVALID_DKEY) {
                    o = rose.grow(op::dedupe::SIZE);
                    wr8(rose.at(o), 0, op::DEDUPE);
                    wr32(rose.at(o), op::dedupe::dkey, f.dkey);
                    wr32(rose.at(o), op::dedupe::fail_jump, (u32)(block_end - o));
                }
                o = rose.grow(op::report_exhaust::SIZE);
                wr8(rose.at(o), 0, op::REPORT_EXHAUST);
                wr32(rose.at(o), op::report_exhaust::onmatch, f.report);
                wr32(rose.at(o), op::report_exhaust::ekey, f.ekey);
            } else if (f.dkey != INVALID_DKEY) {
                size_t o = rose.grow(op::dedupe_and_report::SIZE);
                wr8(rose.at(o), 0, op::DEDUPE_AND_REPORT);
                wr32(rose.at(o), op::dedupe_and_report::dkey, f.dkey);
                wr32(rose.at(o), op::dedupe_and_report::onmatch, f.report);
                wr32(rose.at(o), op::dedupe_and_report::fail_jump, (u32)(block_end - o));
            } else {
                size_t o = rose.grow(op::report::SIZE);
                // a lone trailing REPORT is emitted in its one-shot FINAL_REPORT form
                bool last = bi + 1 == fr.fulls.size();
                wr8(rose.at(o), 0, last ? op::FINAL_REPORT : op::REPORT);
                wr32(rose.at(o), op::report::onmatch, f.report);
            }
            (void)block_start;
        }
        size_t o = rose.grow(op::SIZE_1);
        wr8(rose.at(o), 0, op::END);
    }

    // literal matcher
    std::vector<HwlmLit> hl;
    for (auto &fr : frags) hl.push_back({fr.suffix, fr.nocase, fr.program, 1ull});
    std::vector<u8> engine;
    u8 hwlm_type = hwlm::ENGINE_FDR;
    int engine_choice = opt.engine;
    u32 teddy_sel = 0;
    if (engine_choice == HSB_LIT_ENGINE_AUTO) {
        if (hl.size() == 1) engine_choice = HSB_LIT_ENGINE_NOODLE;
        else if ((teddy_sel = choose_teddy_masks(hl)) != 0) engine_choice = HSB_LIT_ENGINE_TEDDY;
        else engine_choice = HSB_LIT_ENGINE_FDR;
    }
    const bool allow_flood = opt.no_flood == 0;
    if (engine_choice == HSB_LIT_ENGINE_NOODLE) {
        if (hl.size() != 1) { err = "noodle holds exactly one literal fragment"; return HSB_LIT_COMPILER_ERROR; }
        engine = build_noodle(hl[0]);
        hwlm_type = hwlm::ENGINE_NOOD;
    } else if (engine_choice == HSB_LIT_ENGINE_TEDDY) {
        u32 masks = opt.teddy_masks ? (u32)opt.teddy_masks : (teddy_sel & 0xff);
        bool packed = opt.teddy_masks ? hl.size() > 8 : (teddy_sel >> 8) != 0;
        if (!masks) { masks = 1; for (auto &l : hl) masks = std::max<u32>(masks, std::min<u32>(3, (u32)l.s.size())); }
        if (masks < 1 || masks > 4) { err = "teddy_masks must be 1..4"; return HSB_LIT_COMPILER_ERROR; }
        if (hl.size() > 48) { err = "too many literal fragments for Teddy (max 48)"; return HSB_LIT_COMPILER_ERROR; }
        engine = build_teddy(hl, masks, packed, allow_flood, err);
        if (engine.empty()) {
            // as fdrBuildProtoInternal does: "build with teddy failed, will try with FDR"
            if (opt.engine != HSB_LIT_ENGINE_AUTO) return HSB_LIT_COMPILER_ERROR;
            err.clear();
            engine_choice = HSB_LIT_ENGINE_FDR;
        }
    }
    if (engine_choice == HSB_LIT_ENGINE_NOODLE || engine_choice == HSB_LIT_ENGINE_TEDDY) {
        // built above
    } else if (engine_choice == HSB_LIT_ENGINE_FDR) {
        FdrShape shape = choose_fdr_shape(hl);
        if (opt.fdr_domain) shape.domain = (u32)opt.fdr_domain;
        if (opt.fdr_stride) shape.stride = (u32)opt.fdr_stride;
        if (shape.domain < 9 || shape.domain > 15 || (shape.stride != 1 && shape.stride != 2 && shape.stride != 4)) {
            err = "FDR domain must be 9..15 and stride 1, 2 or 4";
            return HSB_LIT_COMPILER_ERROR;
        }
        size_t min_len = ~(size_t)0;
        for (auto &l : hl) min_len = std::min(min_len, l.s.size());
        if (min_len < shape.stride) { err = "FDR stride exceeds the shortest literal"; return HSB_LIT_COMPILER_ERROR; }
        engine = build_fdr(hl, shape, allow_flood);
    } else {
        err = "unknown engine";
        return HSB_LIT_COMPILER_ERROR;
    }

    rose.align(CACHELINE);
    const u32 fmatcher = (u32)rose.grow(hwlm::BODY + engine.size());
    wr8(rose.at(fmatcher), hwlm::type, hwlm_type);
    std::memcpy(rose.at(fmatcher + hwlm::BODY), engine.data(), engine.size());
    rose.align(CACHELINE);

    // engine header
    u32 min_width = ~0u;
    for (auto &f : fulls) min_width = std::min<u32>(min_width, (u32)f.s.size());
    u8 *r = rose.at(0);
    wr8(r, rose::pureLiteral, 1);
    wr8(r, rose::runtimeImpl, rose::RUNTIME_PURE_LITERAL);
    wr8(r, rose::canExhaust, all_exhaustible ? 1 : 0);
    wr32(r, rose::mode, MODE_BLOCK);
    wr32(r, rose::ekeyCount, (u32)ekeys.size());
    wr32(r, rose::dkeyCount, (u32)dkeys.size());
    wr32(r, rose::dkeyLogSize, fatbit_size((u32)dkeys.size()));
    wr32(r, rose::invDkeyOffset, inv_dkey_off);
    wr32(r, rose::somLocationFatbitSize, fatbit_size(0));
    wr32(r, rose::fmatcherOffset, fmatcher);
    wr32(r, rose::fmatcherMinWidth, min_width);
    wr32(r, rose::amatcherMaxBiAnchoredWidth, rose::BOUND_INF);
    wr32(r, rose::fmatcherMaxBiAnchoredWidth, rose::BOUND_INF);
    wr32(r, rose::activeQueueArraySize, fatbit_size(0));
    wr32(r, rose::handledKeyFatbitSize, fatbit_size(0));
    wr32(r, rose::minWidth, min_width);
    wr32(r, rose::minWidthExcludingBoundaries, min_width);
    wr32(r, rose::maxBiAnchoredWidth, rose::BOUND_INF);
    wr32(r, rose::floatingDistance, rose::BOUND_INF);
    wr32(r, rose::smallBlockDistance, 0);
    wr32(r, rose::floatingMinLiteralMatchOffset, min_width);
    wr64(r, rose::initialGroups, 1);
    wr64(r, rose::floating_group_mask, 1);
    wr32(r, rose::size, (u32)rose.size());
    wr32(r, rose::delay_fatbit_size, fatbit_size(0));
    wr32(r, rose::anchored_fatbit_size, fatbit_size(0));
    wr32(r, rose::totalNumLiterals, (u32)frags.size());
    wr32(r, rose::initMpvNfa, 0xffffffffu);
    // state layout (fillStateOffsets): status byte, empty role/leaf/left multibits, 1 group byte,
    // no history, exhaustion multibit
    {
        u8 *so = r + rose::stateOffsets;
        u32 cur = 1;
        wr32(so, rose::so::activeLeafArray, cur);
        wr32(so, rose::so::activeLeftArray, cur);
        wr32(so, rose::so::longLitState, cur);
        wr32(so, rose::so::leftfixLagTable, cur);
        wr32(so, rose::so::anchorState, cur);
        wr32(so, rose::so::groups, cur);
        wr32(so, rose::so::groups_size, 1);
        cur += 1;
        wr32(so, rose::so::history, cur);
        wr32(so, rose::so::exhausted, cur);
        wr32(so, rose::so::exhausted_size, mmbit_size((u32)ekeys.size()));
        cur += mmbit_size((u32)ekeys.size());
        wr32(so, rose::so::logicalVec, cur);
        wr32(so, rose::so::combVec, cur);
        wr32(so, rose::so::nfaStateBegin, cur);
        wr32(so, rose::so::end, cur);
    }

    // database header; the image sits at a 64-byte aligned address with the bytecode at offset 64
    // (inside the header's 64-byte padding field), i.e. cacheline aligned exactly as the reference's
    // db_copy_bytecode arranges (src/database.c:189-197); allocation is sizeof(header) + length
    const size_t bc_off = db::IMAGE_BYTECODE;
    image.assign(db::SIZE + rose.size(), 0);
    std::memcpy(image.data() + bc_off, rose.b.data(), rose.size());
    wr32(image.data(), db::magic, db::MAGIC);
    wr32(image.data(), db::version, db::VERSION_5_4_11);
    wr32(image.data(), db::length, (u32)rose.size());
    wr64(image.data(), db::platform, db::PLATFORM_GENERIC);
    wr32(image.data(), db::crc32, crc32c(rose.b.data(), rose.size()));
    wr32(image.data(), db::bytecode, (u32)bc_off);
    return HSB_LIT_SUCCESS;
}

char *dup_error(const std::string &s) {
    char *p = (char *)std::malloc(s.size() + 1);
    if (p) std::memcpy(p, s.c_str(), s.size() + 1);
    return p;
}

}  // namespace

extern "C" {

int hsb_compile_lit_multi(const char *const *expressions, const unsigned *flags, const unsigned *ids,
                          const size_t *lens, unsigned elements, unsigned mode,
                          const hsb_lit_options *options, hs_database_t **db, char **error) {
    if (error) *error = nullptr;
    if (!db) return HSB_LIT_COMPILER_ERROR;
    *db = nullptr;
    std::string err;
    auto fail = [&](const std::string &m) { if (error) *error = dup_error(m); return HSB_LIT_COMPILER_ERROR; };
    if (!expressions || !lens || !elements) return fail("invalid parameter: no expressions");
    if (mode != hsb::MODE_BLOCK) return fail("only HS_MODE_BLOCK databases are written");
    hsb_lit_options opt{};
    if (options) opt = *options;
    std::vector<FullLit> fulls;
    for (unsigned i = 0; i < elements; i++) {
        unsigned f = flags ? flags[i] : 0;
        if (f & ~(HSB_FLAG_CASELESS | HSB_FLAG_SINGLEMATCH)) return fail("unsupported flag for a pure literal");
        if (!expressions[i] || !lens[i]) return fail("empty literal");
        FullLit fl;
        fl.s.assign(expressions[i], lens[i]);
        fl.nocase = (f & HSB_FLAG_CASELESS) != 0;
        fl.report = ids ? ids[i] : 0;
        fl.single = (f & HSB_FLAG_SINGLEMATCH) != 0;
        fulls.push_back(fl);
    }
    std::vector<u8> image;
    int rv = build(fulls, opt, image, err);
    if (rv != HSB_LIT_SUCCESS) return fail(err);
    void *mem = nullptr;
    if (posix_memalign(&mem, 64, image.size())) return fail("out of memory");
    std::memcpy(mem, image.data(), image.size());
    *db = (hs_database_t *)mem;
    return HSB_LIT_SUCCESS;
}

void hsb_free_database(hs_database_t *db) { std::free(db); }
void hsb_free_error(char *error) { std::free(error); }

int hsb_database_size(const hs_database_t *db, size_t *size) {
    if (!db || !size || hsb::rd32(db, hsb::db::magic) != hsb::db::MAGIC) return HSB_LIT_COMPILER_ERROR;
    *size = hsb::db::SIZE + (size_t)hsb::rd32(db, hsb::db::length);
    return HSB_LIT_SUCCESS;
}

// hs_serialize_database layout (src/database.c:62-110): sizeof(header)+length bytes holding a packed
// 32-byte header {magic, version, length, platform(8), crc32, reserved0, reserved1} then the bytecode
int hsb_serialize_database(const hs_database_t *db, char **bytes, size_t *length) {
    if (!db || !bytes || !length || hsb::rd32(db, hsb::db::magic) != hsb::db::MAGIC) return HSB_LIT_COMPILER_ERROR;
    const u32 len = hsb::rd32(db, hsb::db::length);
    const size_t total = hsb::db::SIZE + len;
    u8 *out = (u8 *)std::calloc(1, total);
    if (!out) return HSB_LIT_COMPILER_ERROR;
    hsb::wr32(out, 0, hsb::rd32(db, hsb::db::magic));
    hsb::wr32(out, 4, hsb::rd32(db, hsb::db::version));
    hsb::wr32(out, 8, len);
    hsb::wr64(out, 12, hsb::rd64(db, hsb::db::platform));
    hsb::wr32(out, 20, hsb::rd32(db, hsb::db::crc32));
    std::memcpy(out + hsb::db::SER_HEADER, (const u8 *)db + hsb::rd32(db, hsb::db::bytecode), len);
    *bytes = (char *)out;
    *length = total;
    return HSB_LIT_SUCCESS;
}

int hsb_deserialize_database(const char *bytes, size_t length, hs_database_t **db) {
    if (!bytes || !db || length < hsb::db::SER_HEADER) return HSB_LIT_COMPILER_ERROR;
    *db = nullptr;
    if (hsb::rd32(bytes, 0) != hsb::db::MAGIC) return HSB_LIT_COMPILER_ERROR;
    const u32 len = hsb::rd32(bytes, 8);
    if (length != hsb::db::SIZE + (size_t)len) return HSB_LIT_COMPILER_ERROR;
    const size_t bc_off = hsb::db::IMAGE_BYTECODE;
    void *mem = nullptr;
    if (posix_memalign(&mem, 64, hsb::db::SIZE + len)) return HSB_LIT_COMPILER_ERROR;
    std::memset(mem, 0, hsb::db::SIZE + len);
    hsb::wr32(mem, hsb::db::magic, hsb::db::MAGIC);
    hsb::wr32(mem, hsb::db::version, hsb::rd32(bytes, 4));
    hsb::wr32(mem, hsb::db::length, len);
    hsb::wr64(mem, hsb::db::platform, hsb::rd64(bytes, 12));
    hsb::wr32(mem, hsb::db::crc32, hsb::rd32(bytes, 20));
    hsb::wr32(mem, hsb::db::bytecode, (u32)bc_off);
    std::memcpy((u8 *)mem + bc_off, bytes + hsb::db::SER_HEADER, len);
    if (crc32c((const u8 *)mem + bc_off, len) != hsb::rd32(mem, hsb::db::crc32)) { std::free(mem); return HSB_LIT_COMPILER_ERROR; }
    *db = (hs_database_t *)mem;
    return HSB_LIT_SUCCESS;
}

unsigned hsb_crc32c(const void *p, size_t n) { return crc32c((const u8 *)p, n); }

}  // extern "C"
