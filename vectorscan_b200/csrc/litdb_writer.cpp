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
#include "outfix_engines.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <map>
#include <set>
#include <string>
#include <tuple>
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

// one byte constraint of a literal's lookaround (offset relative to the literal's end, negative)
struct Look {
    int offset;
    bool is_class;
    u8 andm, cmpm, neg;
    u8 cls[32];
    bool admits(u8 c) const {
        if (is_class) return (cls[c >> 3] >> (c & 7)) & 1;
        return ((c & andm) == cmpm) != (neg != 0);
    }
};

struct FullLit {
    std::string s;
    bool nocase;
    u32 report;
    bool single;  // HS_FLAG_SINGLEMATCH
    u32 ekey = INVALID_EKEY;
    u32 dkey = INVALID_DKEY;
    std::vector<Look> look;  // sorted by offset
    u32 group = 0;
    bool check_groups = false, want_squash = false, quiet = false;
};

struct BuildExt {
    bool included_jump = false, avx512 = false;
    bool full_rose = false;  // runtimeImpl = FULL_ROSE: literal matches run through the full roseRunProgram
    std::vector<hsb_nfadb::Pattern> regexes;  // outfix engines beside the floating literal matcher (implies full_rose)
    unsigned regex_engine = HSB_REGEX_ENGINE_MCCLELLAN;
    u32 dfa_states = 0;
    u64 groups_off = 0;
    u32 max_bi_anchored_width = 0;
    bool vectored = false;  // HS_MODE_VECTORED: the buffers of one hs_scan_vector call are consecutive blocks of ONE stream
    bool streaming = false; // HS_MODE_STREAM: the same stream state, driven through hs_open_stream / hs_scan_stream
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

// nbuckets = 8: Teddy (engines 11-18); nbuckets = 16: fat Teddy (engines 3-10, AVX2 hosts): every nibble mask is laid out
// twice, buckets 0-7 then buckets 8-15 (maskWidth 2, teddy_compile.cpp:436-500), and the reinforcement table's place is
// taken by the "dup" masks -- each 16-byte mask twice in a row (fillDupNibbleMasks :355-435; build :559-620).
std::vector<u8> build_teddy(const std::vector<HwlmLit> &lits, u32 masks, bool packed, bool allow_flood,
                            std::string &err, u32 nbuckets = 8) {
    std::vector<std::vector<u32>> buckets;
    if (!pack_teddy(lits, masks, nbuckets, buckets)) { err = "literal set does not pack into the Teddy buckets"; return {}; }
    const u32 width = nbuckets / 8;
    const size_t mask_len = (size_t)masks * 32 * width;
    const size_t rtab_len = width == 1 ? teddy::REINFORCE_ROWS * 8 : mask_len * 2;
    std::vector<u8> nib(mask_len, 0xff), dup(width == 2 ? mask_len * 2 : 0, 0xff);
    std::vector<u64> reinforce(teddy::REINFORCE_ROWS, 0x00ffffffffffffffull);
    for (u32 b = 0; b < nbuckets; b++) {
        const u8 bm = (u8)(1u << (b % 8));
        const u32 g = b / 8;
        for (u32 li : buckets[b]) {
            const HwlmLit &l = lits[li];
            const u32 sz = (u32)l.s.size();
            for (u32 j = 0; j < masks; j++) {
                const size_t lo_id = (size_t)(2 * j) * width + g, hi_id = (size_t)(2 * j + 1) * width + g;
                u8 *lo = nib.data() + lo_id * 16, *hi = nib.data() + hi_id * 16;
                auto clear_lo = [&](u32 n) { lo[n] &= (u8)~bm; if (width == 2) { dup[lo_id * 32 + n] &= (u8)~bm; dup[lo_id * 32 + 16 + n] &= (u8)~bm; } };
                auto clear_hi = [&](u32 n) { hi[n] &= (u8)~bm; if (width == 2) { dup[hi_id * 32 + n] &= (u8)~bm; dup[hi_id * 32 + 16 + n] &= (u8)~bm; } };
                if (j >= sz) {
                    for (u32 n = 0; n < 16; n++) { clear_lo(n); clear_hi(n); }
                    continue;
                }
                u8 c = (u8)l.s[sz - 1 - j];
                clear_lo(c & 0xf);
                if (l.nocase && is_alpha(c)) {
                    clear_hi((c >> 4) & 0xd);
                    clear_hi((c >> 4) | 0x2);
                } else {
                    clear_hi(c >> 4);
                }
            }
            if (width != 1) continue;
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
    wr32(out.data(), fdr::engineID, (width == 2 ? teddy::FAT_FIRST : teddy::ENGINE_FIRST) + 2 * (masks - 1) + (packed ? 1 : 0));
    wr32(out.data(), fdr::size, (u32)total);
    wr32(out.data(), fdr::maxStringLen, max_len);
    wr32(out.data(), fdr::numStrings, (u32)lits.size());
    wr32(out.data(), fdr::confOffset, (u32)conf_off);
    wr32(out.data(), fdr::floodOffset, (u32)flood_off);
    std::memcpy(out.data() + teddy::MASKS, nib.data(), mask_len);
    if (width == 1) std::memcpy(out.data() + teddy::MASKS + round_up(mask_len, CACHELINE), reinforce.data(), rtab_len);
    else std::memcpy(out.data() + teddy::MASKS + round_up(mask_len, CACHELINE), dup.data(), rtab_len);
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
// The instruction that checks a literal's lookaround `look` (sorted by offset), chosen as makeLookaroundInstruction
// does (src/rose/rose_build_program.cpp:1343-1523): and/cmp entries -> CHECK_BYTE / CHECK_MASK / CHECK_MASK_32 /
// CHECK_MASK_64 (makeRoleByte :969, makeRoleMask :989, makeRoleMask32 :1034, makeRoleMask64 :1074); class entries ->
// the shufti family (makeRoleShufti :1380, mask layout getShuftiMasks :1190 / nibMaskUpdate :1175).  `fails` receives
// (instruction offset, offset of its fail_jump operand) so the caller can aim it at the end of the block.
bool emit_look(Buf &rose, const std::vector<Look> &look, bool avx512, bool full_rose,
               std::vector<std::pair<size_t, size_t>> &fails, std::string &err) {
    if (look.empty()) return true;
    const int base = look.front().offset, span = look.back().offset - base + 1;
    bool any_class = false;
    for (auto &e : look) any_class |= e.is_class;
    if (!any_class) {
        if (look.size() == 1) {
            const Look &e = look[0];
            size_t o = rose.grow(op::check_byte::SIZE);
            wr8(rose.at(o), 0, op::CHECK_BYTE);
            wr8(rose.at(o), op::check_byte::and_mask, e.andm);
            wr8(rose.at(o), op::check_byte::cmp_mask, e.cmpm);
            wr8(rose.at(o), op::check_byte::negation, e.neg ? 1 : 0);
            wr32(rose.at(o), op::check_byte::offset, (u32)e.offset);
            fails.push_back({o, op::check_byte::fail_jump});
            return true;
        }
        if (span <= 8) {
            u64 am = 0, cm = 0, nm = 0;
            for (auto &e : look) {
                const unsigned sh = 8u * (unsigned)(e.offset - base);
                am |= (u64)e.andm << sh;
                cm |= (u64)e.cmpm << sh;
                if (e.neg) nm |= 0xffull << sh;
            }
            size_t o = rose.grow(op::check_mask::SIZE);
            wr8(rose.at(o), 0, op::CHECK_MASK);
            wr64(rose.at(o), op::check_mask::and_mask, am);
            wr64(rose.at(o), op::check_mask::cmp_mask, cm);
            wr64(rose.at(o), op::check_mask::neg_mask, nm);
            wr32(rose.at(o), op::check_mask::offset, (u32)base);
            fails.push_back({o, op::check_mask::fail_jump});
            return true;
        }
        if (span <= 32 || (span <= 64 && avx512)) {
            const bool wide = span > 32;
            size_t o = rose.grow(wide ? op::check_mask64::SIZE : op::check_mask32::SIZE);
            wr8(rose.at(o), 0, wide ? op::CHECK_MASK_64 : op::CHECK_MASK_32);
            u64 nm = 0;
            for (auto &e : look) {
                const size_t k = (size_t)(e.offset - base);
                wr8(rose.at(o), (wide ? op::check_mask64::and_mask : op::check_mask32::and_mask) + k, e.andm);
                wr8(rose.at(o), (wide ? op::check_mask64::cmp_mask : op::check_mask32::cmp_mask) + k, e.cmpm);
                if (e.neg) nm |= 1ull << k;
            }
            if (wide) {
                wr64(rose.at(o), op::check_mask64::neg_mask, nm);
                wr32(rose.at(o), op::check_mask64::offset, (u32)base);
                fails.push_back({o, op::check_mask64::fail_jump});
            } else {
                wr32(rose.at(o), op::check_mask32::neg_mask, (u32)nm);
                wr32(rose.at(o), op::check_mask32::offset, (u32)base);
                fails.push_back({o, op::check_mask32::fail_jump});
            }
            return true;
        }
        err = "lookaround spans more bytes than any mask instruction of the target holds";
        return false;
    }
    // roseRunProgram_l, the interpreter of pure-literal databases, has no CHECK_SHUFTI_* case (program_runtime.c:3083-3504)
    if (!full_rose) { err = "class lookaround needs a full-Rose database (the pure-literal interpreter has no CHECK_SHUFTI)"; return false; }
    // shufti family: a bucket = one (high-nibble set) x (low-nibble set) product; a class is the union of its products
    if (span > (avx512 ? 64 : 32)) { err = "class lookaround spans more bytes than a shufti instruction holds"; return false; }
    std::map<u32, u32> bucket_of;  // (hi set << 16 | lo set) -> bucket index
    u8 hi_mask[32] = {0}, lo_mask[32] = {0}, sel_lo[64] = {0}, sel_hi[64] = {0};
    u64 neg = ~0ull;
    for (auto &e : look) {
        bool in[256];
        u32 count = 0;
        for (u32 c = 0; c < 256; c++) { in[c] = e.admits((u8)c); count += in[c]; }
        const size_t k = (size_t)(e.offset - base);
        if (count > 128) { for (u32 c = 0; c < 256; c++) in[c] = !in[c]; }  // heavy classes are stored complemented
        else neg ^= 1ull << k;
        std::map<u32, u32> lo2hi;
        for (u32 h = 0; h < 16; h++) {
            u32 lo = 0;
            for (u32 l = 0; l < 16; l++) if (in[h * 16 + l]) lo |= 1u << l;
            if (lo) lo2hi[lo] |= 1u << h;
        }
        for (auto &kv : lo2hi) {
            const u32 key = (kv.second << 16) | kv.first;
            auto it = bucket_of.find(key);
            if (it == bucket_of.end()) {
                const u32 b = (u32)bucket_of.size();
                if (b >= 16) { err = "class lookaround needs more than 16 shufti buckets"; return false; }
                it = bucket_of.emplace(key, b).first;
                for (u32 n = 0; n < 16; n++) {
                    if ((kv.second >> n) & 1) hi_mask[n + (b >= 8 ? 16 : 0)] |= (u8)(1u << (b & 7));
                    if ((kv.first >> n) & 1) lo_mask[n + (b >= 8 ? 16 : 0)] |= (u8)(1u << (b & 7));
                }
            }
            if (it->second >= 8) sel_hi[k] |= (u8)(1u << (it->second - 8)); else sel_lo[k] |= (u8)(1u << it->second);
        }
    }
    const u32 nb = (u32)bucket_of.size();
    auto rep4 = [&](size_t o, size_t field, const u8 *m16) { for (int r = 0; r < 4; r++) std::memcpy(rose.at(o) + field + 16 * r, m16, 16); };
    if (span <= 16 && nb <= 8) {
        namespace I = op::check_shufti_16x8;
        size_t o = rose.grow(I::SIZE);
        wr8(rose.at(o), 0, op::CHECK_SHUFTI_16x8);
        std::memcpy(rose.at(o) + I::nib_mask, lo_mask, 16);
        std::memcpy(rose.at(o) + I::nib_mask + 16, hi_mask, 16);
        std::memcpy(rose.at(o) + I::bucket_select_mask, sel_lo, 16);
        wr32(rose.at(o), I::neg_mask, (u32)(neg & 0xffff));
        wr32(rose.at(o), I::offset, (u32)base);
        fails.push_back({o, I::fail_jump});
    } else if (span <= 32 && nb <= 8) {
        namespace I = op::check_shufti_32x8;
        size_t o = rose.grow(I::SIZE);
        wr8(rose.at(o), 0, op::CHECK_SHUFTI_32x8);
        std::memcpy(rose.at(o) + I::hi_mask, hi_mask, 16);
        std::memcpy(rose.at(o) + I::lo_mask, lo_mask, 16);
        std::memcpy(rose.at(o) + I::bucket_select_mask, sel_lo, 32);
        wr32(rose.at(o), I::neg_mask, (u32)neg);
        wr32(rose.at(o), I::offset, (u32)base);
        fails.push_back({o, I::fail_jump});
    } else if (avx512 && nb <= 8) {
        namespace I = op::check_shufti_64x8;
        size_t o = rose.grow(I::SIZE);
        wr8(rose.at(o), 0, op::CHECK_SHUFTI_64x8);
        rep4(o, I::hi_mask, hi_mask);
        rep4(o, I::lo_mask, lo_mask);
        std::memcpy(rose.at(o) + I::bucket_select_mask, sel_lo, 64);
        wr64(rose.at(o), I::neg_mask, neg);
        wr32(rose.at(o), I::offset, (u32)base);
        fails.push_back({o, I::fail_jump});
    } else if (span <= 16) {
        namespace I = op::check_shufti_16x16;
        size_t o = rose.grow(I::SIZE);
        wr8(rose.at(o), 0, op::CHECK_SHUFTI_16x16);
        std::memcpy(rose.at(o) + I::hi_mask, hi_mask, 32);
        std::memcpy(rose.at(o) + I::lo_mask, lo_mask, 32);
        std::memcpy(rose.at(o) + I::bucket_select_mask, sel_lo, 16);
        std::memcpy(rose.at(o) + I::bucket_select_mask + 16, sel_hi, 16);
        wr32(rose.at(o), I::neg_mask, (u32)(neg & 0xffff));
        wr32(rose.at(o), I::offset, (u32)base);
        fails.push_back({o, I::fail_jump});
    } else if (span <= 32) {
        namespace I = op::check_shufti_32x16;
        size_t o = rose.grow(I::SIZE);
        wr8(rose.at(o), 0, op::CHECK_SHUFTI_32x16);
        std::memcpy(rose.at(o) + I::hi_mask, hi_mask, 32);
        std::memcpy(rose.at(o) + I::lo_mask, lo_mask, 32);
        std::memcpy(rose.at(o) + I::bucket_select_mask_hi, sel_hi, 32);
        std::memcpy(rose.at(o) + I::bucket_select_mask_lo, sel_lo, 32);
        wr32(rose.at(o), I::neg_mask, (u32)neg);
        wr32(rose.at(o), I::offset, (u32)base);
        fails.push_back({o, I::fail_jump});
    } else {
        namespace I = op::check_shufti_64x16;
        size_t o = rose.grow(I::SIZE);
        wr8(rose.at(o), 0, op::CHECK_SHUFTI_64x16);
        rep4(o, I::hi_mask_1, hi_mask);
        rep4(o, I::hi_mask_2, hi_mask + 16);
        rep4(o, I::lo_mask_1, lo_mask);
        rep4(o, I::lo_mask_2, lo_mask + 16);
        std::memcpy(rose.at(o) + I::bucket_select_mask_hi, sel_hi, 64);
        std::memcpy(rose.at(o) + I::bucket_select_mask_lo, sel_lo, 64);
        wr64(rose.at(o), I::neg_mask, neg);
        wr32(rose.at(o), I::offset, (u32)base);
        fails.push_back({o, I::fail_jump});
    }
    return true;
}

struct Fragment {
    std::string suffix;  // <= 8 bytes handed to the literal matcher
    bool nocase;
    std::vector<u32> fulls;  // indices into the FullLit array
    u32 program = 0;
    u32 group = 0;
    int child = -1;          // fragment whose text is a suffix of this one's and whose program INCLUDED_JUMP enters
    u8 squash = 0;           // FDR confirm bits (buckets) of that fragment
};

int build(const std::vector<FullLit> &fulls_in, const hsb_lit_options &opt, const BuildExt &bx, std::vector<u8> &image,
          std::string &err) {
    std::vector<FullLit> fulls = fulls_in;
    if (fulls.empty()) { err = "no literals"; return HSB_LIT_COMPILER_ERROR; }
    for (auto &f : fulls)
        if (f.s.empty()) { err = "empty literal"; return HSB_LIT_COMPILER_ERROR; }

    const bool full_rose = bx.full_rose || !bx.regexes.empty();
    for (auto &p : bx.regexes)
        for (auto &f : fulls)
            if (f.report == p.id) { err = "a report id belongs either to literals or to regexes"; return HSB_LIT_COMPILER_ERROR; }
    // exhaustion keys: one per SINGLEMATCH report id; dedupe keys: report ids reachable from more
    // than one distinct literal (the same (id, to) could otherwise be delivered twice)
    std::map<u32, u32> ekeys, dkeys;
    std::map<u32, std::set<std::pair<std::string, bool>>> lits_per_report;
    std::map<u32, int> quiet_of;  // report id -> quiet? (one report = one behaviour)
    for (auto &f : fulls) {
        if (f.single) ekeys.emplace(f.report, 0);
        lits_per_report[f.report].insert({f.s, f.nocase});
        auto q = quiet_of.emplace(f.report, f.quiet ? 1 : 0);
        if (q.first->second != (f.quiet ? 1 : 0)) { err = "a report id is either quiet or not"; return HSB_LIT_COMPILER_ERROR; }
        if (f.group > 62) { err = "group must be 0..62"; return HSB_LIT_COMPILER_ERROR; }
        for (size_t k = 0; k < f.look.size(); k++)
            if (f.look[k].offset < -64 || f.look[k].offset > -1 || (k && f.look[k].offset <= f.look[k - 1].offset)) {
                err = "lookaround offsets must be distinct, ascending and within -64..-1";
                return HSB_LIT_COMPILER_ERROR;
            }
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
    std::map<std::tuple<std::string, bool, u32>, u32> frag_index;
    std::vector<Fragment> frags;
    for (u32 i = 0; i < fulls.size(); i++) {
        const FullLit &f = fulls[i];
        std::string suf = f.s.size() > 8 ? f.s.substr(f.s.size() - 8) : f.s;
        bool any_alpha = false;
        for (char c : suf) any_alpha |= is_alpha((u8)c);
        auto key = std::make_tuple(suf, f.nocase && any_alpha, f.group);
        auto it = frag_index.find(key);
        if (it == frag_index.end()) {
            it = frag_index.emplace(key, (u32)frags.size()).first;
            Fragment fr;
            fr.suffix = suf;
            fr.nocase = f.nocase && any_alpha;
            fr.group = f.group;
            frags.push_back(fr);
        }
        // identical (text, case, report) duplicates add nothing
        bool dup = false;
        for (u32 o : frags[it->second].fulls)
            dup |= fulls[o].s == f.s && fulls[o].nocase == f.nocase && fulls[o].report == f.report &&
                   fulls[o].look.empty() && f.look.empty();
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

    // Suffix inclusion (FDR only; findIncludedLits / buildSquashMask, src/fdr/fdr_compile.cpp:654-812): when fragment P
    // confirms at some end position and fragment C -- a proper suffix of P with the same caselessness and group, alone
    // among the fragments of its bucket that can end where P ends -- sits in a LATER bucket, P's program ends in an
    // INCLUDED_JUMP that clears C's bucket bit for that position and runs C's program directly.  This writer only
    // takes the clear-cut cases (one child per parent, one parent per child); leaving a squash out is always valid.
    if (bx.included_jump) {
        if (opt.engine != HSB_LIT_ENGINE_FDR) { err = "included_jump needs engine = FDR"; return HSB_LIT_COMPILER_ERROR; }
        std::vector<HwlmLit> pre;
        for (u32 i = 0; i < frags.size(); i++) pre.push_back({frags[i].suffix, frags[i].nocase, i, 1ull << frags[i].group});
        std::vector<std::vector<u32>> pb = assign_fdr_buckets(pre);
        std::vector<u32> bucket_of(frags.size(), 0);
        for (u32 b = 0; b < pb.size(); b++) for (u32 li : pb[b]) bucket_of[pre[li].id] = b;
        auto upper = [](const std::string &x) { std::string r = x; for (auto &c : r) c = (char)to_upper((u8)c); return r; };
        auto ends_with = [](const std::string &a, const std::string &b) {
            return b.size() <= a.size() && a.compare(a.size() - b.size(), b.size(), b) == 0;
        };
        std::vector<bool> has_parent(frags.size(), false);
        for (u32 p = 0; p < frags.size(); p++) {
            const std::string up = upper(frags[p].suffix);
            bool same_bucket = false;
            for (u32 x = 0; x < frags.size(); x++)
                same_bucket |= x != p && bucket_of[x] == bucket_of[p] && ends_with(up, upper(frags[x].suffix));
            if (same_bucket) continue;
            for (u32 b2 = bucket_of[p] + 1; b2 < fdr::NUM_BUCKETS; b2++) {
                std::vector<u32> related;
                for (u32 x = 0; x < frags.size(); x++)
                    if (bucket_of[x] == b2 && ends_with(up, upper(frags[x].suffix))) related.push_back(x);
                if (related.empty()) continue;
                if (related.size() == 1) {
                    const u32 c = related[0];
                    const Fragment &P = frags[p], &Cf = frags[c];
                    const bool same_text = P.nocase ? true : ends_with(P.suffix, Cf.suffix);
                    if (Cf.suffix.size() < P.suffix.size() && Cf.nocase == P.nocase && Cf.group == P.group && same_text &&
                        !has_parent[c]) {
                        frags[p].child = (int)c;
                        frags[p].squash = (u8)(1u << b2);
                        has_parent[c] = true;
                    }
                }
                break;  // only the first later bucket holding a suffix is considered
            }
        }
    }
    std::map<u32, u32> group_users;
    for (auto &fr : frags) group_users[fr.group]++;

    // programs: [CHECK_GROUPS] then per full literal a block  [lookaround] [CHECK_MED_LIT*] [CHECK_EXHAUSTED] [DEDUPE]
    // REPORT*  whose fail_jumps land on the next block, then [SQUASH_GROUPS] [INCLUDED_JUMP]; END closes the program
    std::vector<std::pair<size_t, u32>> jump_patches;  // INCLUDED_JUMP instruction -> child fragment
    for (auto &fr : frags) {
        rose.align(op::ALIGN);
        fr.program = (u32)rose.size();
        if (!bx.regexes.empty()) {
            // outfix engines must deliver their earlier matches first (makeCatchup, rose_build_program.cpp; CATCH_UP ->
            // roseCatchUpTo, program_runtime.c:2499-2504)
            size_t o = rose.grow(op::SIZE_1);
            wr8(rose.at(o), 0, op::CATCH_UP);
        }
        bool any_check_groups = false;
        for (u32 i : fr.fulls) any_check_groups |= fulls[i].check_groups;
        if (any_check_groups) {
            size_t o = rose.grow(op::check_groups::SIZE);
            wr8(rose.at(o), 0, op::CHECK_GROUPS);
            wr64(rose.at(o), op::check_groups::groups, 1ull << fr.group);
        }
        const bool squash = fr.fulls.size() == 1 && fulls[fr.fulls[0]].want_squash && fulls[fr.fulls[0]].single &&
                            group_users[fr.group] == 1;
        const bool trailing = squash || fr.child >= 0;  // (FINAL_REPORT only closes a program that has nothing behind it)
        for (size_t bi = 0; bi < fr.fulls.size(); bi++) {
            const FullLit &f = fulls[fr.fulls[bi]];
            std::vector<std::pair<size_t, size_t>> fails;
            if (!emit_look(rose, f.look, bx.avx512, full_rose, fails, err)) return HSB_LIT_COMPILER_ERROR;
            if (f.s.size() > 8) {
                size_t o = rose.grow(op::check_lit::SIZE);
                wr8(rose.at(o), 0, f.nocase ? op::CHECK_MED_LIT_NOCASE : op::CHECK_MED_LIT);
                wr32(rose.at(o), op::check_lit::lit_offset, lit_text_off[fr.fulls[bi]]);
                wr32(rose.at(o), op::check_lit::lit_length, (u32)f.s.size());
                fails.push_back({o, op::check_lit::fail_jump});
            }
            if (f.ekey != INVALID_EKEY) {
                size_t o = rose.grow(op::check_exhausted::SIZE);
                wr8(rose.at(o), 0, op::CHECK_EXHAUSTED);
                wr32(rose.at(o), op::check_exhausted::ekey, f.ekey);
                fails.push_back({o, op::check_exhausted::fail_jump});
                if (f.dkey != INVALID_DKEY) {
                    o = rose.grow(op::dedupe::SIZE);
                    wr8(rose.at(o), 0, op::DEDUPE);
                    wr32(rose.at(o), op::dedupe::dkey, f.dkey);
                    fails.push_back({o, op::dedupe::fail_jump});
                }
                if (f.quiet) {
                    o = rose.grow(op::set_exhaust::SIZE);
                    wr8(rose.at(o), 0, op::SET_EXHAUST);
                    wr32(rose.at(o), op::set_exhaust::ekey, f.ekey);
                } else {
                    o = rose.grow(op::report_exhaust::SIZE);
                    wr8(rose.at(o), 0, op::REPORT_EXHAUST);
                    wr32(rose.at(o), op::report_exhaust::onmatch, f.report);
                    wr32(rose.at(o), op::report_exhaust::ekey, f.ekey);
                }
            } else if (f.quiet) {
                // a quiet report without an exhaustion key delivers nothing
            } else if (f.dkey != INVALID_DKEY) {
                size_t o = rose.grow(op::dedupe_and_report::SIZE);
                wr8(rose.at(o), 0, op::DEDUPE_AND_REPORT);
                wr32(rose.at(o), op::dedupe_and_report::dkey, f.dkey);
                wr32(rose.at(o), op::dedupe_and_report::onmatch, f.report);
                fails.push_back({o, op::dedupe_and_report::fail_jump});
            } else {
                size_t o = rose.grow(op::report::SIZE);
                // a lone trailing REPORT is emitted in its one-shot FINAL_REPORT form
                bool last = bi + 1 == fr.fulls.size() && !trailing;
                wr8(rose.at(o), 0, last ? op::FINAL_REPORT : op::REPORT);
                wr32(rose.at(o), op::report::onmatch, f.report);
            }
            const size_t block_end = rose.size();
            for (auto &fj : fails) wr32(rose.at(fj.first), fj.second, (u32)(block_end - fj.first));
        }
        if (squash) {
            size_t o = rose.grow(op::squash_groups::SIZE);
            wr8(rose.at(o), 0, op::SQUASH_GROUPS);
            wr64(rose.at(o), op::squash_groups::groups, ~(1ull << fr.group));
        }
        if (fr.child >= 0) {
            size_t o = rose.grow(op::included_jump::SIZE);
            wr8(rose.at(o), 0, op::INCLUDED_JUMP);
            wr8(rose.at(o), op::included_jump::squash, fr.squash);
            jump_patches.push_back({o, (u32)fr.child});
        }
        size_t o = rose.grow(op::SIZE_1);
        wr8(rose.at(o), 0, op::END);
    }
    for (auto &jp : jump_patches) wr32(rose.at(jp.first), op::included_jump::child_offset, frags[jp.second].program);

    // literal matcher
    std::vector<HwlmLit> hl;
    for (auto &fr : frags) hl.push_back({fr.suffix, fr.nocase, fr.program, 1ull << fr.group});
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
        bool packed = opt.teddy_masks ? hl.size() > (opt.teddy_fat ? 16u : 8u) : (teddy_sel >> 8) != 0;
        if (!masks) { masks = 1; for (auto &l : hl) masks = std::max<u32>(masks, std::min<u32>(3, (u32)l.s.size())); }
        if (masks < 1 || masks > 4) { err = "teddy_masks must be 1..4"; return HSB_LIT_COMPILER_ERROR; }
        const u32 tb = opt.teddy_fat ? 16u : 8u;
        if (opt.teddy_fat && !opt.teddy_masks) packed = hl.size() > 16;
        if (hl.size() > 6 * tb) { err = "too many literal fragments for Teddy (max 6 per bucket)"; return HSB_LIT_COMPILER_ERROR; }
        engine = build_teddy(hl, masks, packed, allow_flood, err, tb);
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
    const u32 fmatcher_min_width = min_width;

    // Outfix engines beside the literal matcher (mixed literal + regex sets): report programs, one engine per queue,
    // NfaInfo array, state_init plan -- the layout of nfadb_writer.cpp build_rose.
    u32 n_queues = 0, nfa_info_off = 0, leaf_bytes = 0, scratch_total = 0, total_dkeys = (u32)dkeys.size();
    u32 plan_u64_off = 0, plan_u64_n = 0, plan_u8_off = 0, plan_u8_n = 0, eod_program = 0;
    std::vector<hsb_nfadb::EngineBlob> blobs;
    std::vector<u32> nfa_at;
    const u32 leaf_off = 1;
    if (!bx.regexes.empty()) {
        std::map<u32, u32> rdkeys;  // regex report id -> dedupe key, numbered behind the literal ones
        for (auto &p : bx.regexes) if (!rdkeys.count(p.id)) { const u32 k = total_dkeys++; rdkeys[p.id] = k; }
        rose.align(8);
        inv_dkey_off = (u32)rose.grow(4 * (size_t)total_dkeys);
        for (auto &d : dkeys) wr32(rose.at(inv_dkey_off), 4 * d.second, d.first);
        for (auto &d : rdkeys) wr32(rose.at(inv_dkey_off), 4 * d.second, d.first);
        std::vector<u32> program_of(bx.regexes.size());
        for (size_t pi = 0; pi < bx.regexes.size(); pi++) {
            rose.align(op::ALIGN);
            const u32 o = (u32)rose.grow(op::dedupe_and_report::SIZE + op::SIZE_1);
            program_of[pi] = o;
            wr8(rose.at(o), 0, op::DEDUPE_AND_REPORT);
            wr32(rose.at(o), op::dedupe_and_report::dkey, rdkeys[bx.regexes[pi].id]);
            wr32(rose.at(o), op::dedupe_and_report::onmatch, bx.regexes[pi].id);
            wr32(rose.at(o), op::dedupe_and_report::fail_jump, (u32)op::dedupe_and_report::SIZE);
            wr8(rose.at(o + op::dedupe_and_report::SIZE), 0, op::END);
        }
        if (hsb_nfadb::build_outfix_engines(bx.regexes, bx.regex_engine, bx.dfa_states, program_of, blobs, err) != HSB_LIT_SUCCESS)
            return HSB_LIT_COMPILER_ERROR;
        n_queues = (u32)blobs.size();
        leaf_bytes = mmbit_size(n_queues);
        rose.align(4);
        nfa_info_off = (u32)rose.grow((size_t)nfainfo::SIZE * n_queues);
        for (u32 qi = 0; qi < n_queues; qi++) {
            rose.align(CACHELINE);
            const u32 at = (u32)rose.grow(blobs[qi].bytes.size());
            std::memcpy(rose.at(at), blobs[qi].bytes.data(), blobs[qi].bytes.size());
            nfa_at.push_back(at);
            min_width = std::min(min_width, blobs[qi].min_width);
        }
        eod_program = hsb_nfadb::emit_engines_eod(rose, blobs, err);
        if (eod_program == ~0u) return HSB_LIT_COMPILER_ERROR;
        for (u32 qi = 0; qi < n_queues; qi++)
            wr8(rose.at(nfa_info_off + (size_t)nfainfo::SIZE * qi), nfainfo::eod, hsb_nfadb::accepts_eod(blobs[qi]) ? 1 : 0);
        rose.align(CACHELINE);
        u32 b = 0;
        if (n_queues > 256) {  // hierarchical multibit: clearing the root block empties it (src/util/multibit.h:332-343)
            plan_u64_off = (u32)rose.size();
            wr32(rose.at(rose.grow(16)), 0, leaf_off);
            plan_u64_n = 1;
            b = leaf_bytes;
        } else if (leaf_bytes >= 8) {
            plan_u64_off = (u32)rose.size();
            for (; b + 8 <= leaf_bytes; b += 8) { wr32(rose.at(rose.grow(16)), 0, leaf_off + b); plan_u64_n++; }
        }
        if (b < leaf_bytes) {
            plan_u8_off = (u32)rose.size();
            for (; b < leaf_bytes; b++) { wr32(rose.at(rose.grow(8)), 0, leaf_off + b); plan_u8_n++; }
        }
        rose.align(CACHELINE);
    }

    u8 *r = rose.at(0);
    wr8(r, rose::pureLiteral, full_rose ? 0 : 1);
    wr8(r, rose::runtimeImpl, full_rose ? rose::RUNTIME_FULL_ROSE : rose::RUNTIME_PURE_LITERAL);
    wr8(r, rose::hasOutfixesInSmallBlock, n_queues ? 1 : 0);
    wr8(r, rose::requiresEodCheck, eod_program ? 1 : 0);
    wr32(r, rose::eodProgramOffset, eod_program);
    wr8(r, rose::canExhaust, all_exhaustible && bx.regexes.empty() ? 1 : 0);
    wr32(r, rose::mode, bx.vectored ? MODE_VECTORED : bx.streaming ? MODE_STREAM : MODE_BLOCK);
    // Stream state keeps the last historyRequired bytes of the stream (src/runtime.c:77-85, maintainHistoryBuffer): the
    // literal confirm (fdr_confirm_runtime.h:77-88) and CHECK_MED_LIT (program_runtime.c:1988-2011) read a literal's
    // bytes that lie before the current block from there, so it must hold longest literal - 1 bytes.
    u32 history = 0;
    if (bx.vectored || bx.streaming) {
        for (auto &f : fulls_in) {
            history = std::max<u32>(history, (u32)f.s.size() ? (u32)f.s.size() - 1 : 0);
            for (auto &l : f.look)   // lookaround bytes before the block come from the history too (program_runtime.c:1253-1340)
                if (l.offset < 0) history = std::max<u32>(history, (u32)(-l.offset));
        }
    }
    wr32(r, rose::historyRequired, history);
    wr32(r, rose::ekeyCount, (u32)ekeys.size());
    wr32(r, rose::dkeyCount, total_dkeys);
    wr32(r, rose::dkeyLogSize, fatbit_size(total_dkeys));
    wr32(r, rose::invDkeyOffset, inv_dkey_off);
    wr32(r, rose::somLocationFatbitSize, fatbit_size(0));
    wr32(r, rose::fmatcherOffset, fmatcher);
    wr32(r, rose::fmatcherMinWidth, fmatcher_min_width);
    wr32(r, rose::amatcherMaxBiAnchoredWidth, rose::BOUND_INF);
    wr32(r, rose::fmatcherMaxBiAnchoredWidth, rose::BOUND_INF);
    wr32(r, rose::activeArrayCount, n_queues);
    wr32(r, rose::queueCount, n_queues);
    wr32(r, rose::activeQueueArraySize, fatbit_size(n_queues));
    wr32(r, rose::nfaInfoOffset, nfa_info_off);
    wr32(r, rose::outfixBeginQueue, 0);
    wr32(r, rose::outfixEndQueue, n_queues);
    wr32(r, rose::leftfixBeginQueue, n_queues);
    wr32(r, rose::handledKeyFatbitSize, fatbit_size(0));
    wr32(r, rose::minWidth, min_width);
    wr32(r, rose::minWidthExcludingBoundaries, min_width);
    wr32(r, rose::maxBiAnchoredWidth, bx.max_bi_anchored_width ? bx.max_bi_anchored_width : rose::BOUND_INF);
    wr32(r, rose::floatingDistance, rose::BOUND_INF);
    wr32(r, rose::smallBlockDistance, 0);
    wr32(r, rose::floatingMinLiteralMatchOffset, full_rose ? 0 : fmatcher_min_width);
    u64 all_groups = 0;
    u32 top_group = 0;
    for (auto &fr : frags) { all_groups |= 1ull << fr.group; top_group = std::max(top_group, fr.group); }
    wr64(r, rose::initialGroups, all_groups & ~bx.groups_off);
    wr64(r, rose::floating_group_mask, all_groups);
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
        wr32(so, rose::so::activeLeafArray_size, leaf_bytes);
        cur += leaf_bytes;
        wr32(so, rose::so::activeLeftArray, cur);
        wr32(so, rose::so::longLitState, cur);
        wr32(so, rose::so::leftfixLagTable, cur);
        wr32(so, rose::so::anchorState, cur);
        wr32(so, rose::so::groups, cur);
        wr32(so, rose::so::groups_size, top_group / 8 + 1);
        cur += top_group / 8 + 1;
        wr32(so, rose::so::history, cur);
        cur += history;
        wr32(so, rose::so::exhausted, cur);
        wr32(so, rose::so::exhausted_size, mmbit_size((u32)ekeys.size()));
        cur += mmbit_size((u32)ekeys.size());
        wr32(so, rose::so::logicalVec, cur);
        wr32(so, rose::so::combVec, cur);
        wr32(so, rose::so::nfaStateBegin, cur);
        for (u32 qi = 0; qi < n_queues; qi++) {
            u8 *ni = rose.at(nfa_info_off + (size_t)nfainfo::SIZE * qi);
            wr32(ni, nfainfo::nfaOffset, nfa_at[qi]);
            wr32(ni, nfainfo::stateOffset, cur);
            scratch_total = (u32)round_up(scratch_total, 64);
            wr32(ni, nfainfo::fullStateOffset, scratch_total);
            cur += blobs[qi].stream_state;
            scratch_total += blobs[qi].scratch_state;
        }
        wr32(so, rose::so::end, cur);
        wr32(r, rose::stateSize, cur);
        wr32(r, rose::scratchStateSize, (u32)round_up(scratch_total, 64));
        u8 *sp = r + rose::state_init;  // struct scatter_full_plan (src/util/scatter.h:43-52)
        wr32(sp, 0, plan_u64_off); wr32(sp, 4, plan_u64_n);
        wr32(sp, 24, plan_u8_n); wr32(sp, 28, plan_u8_off);
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

int hsb_compile_mixed(const char *const *expressions, const unsigned *flags, const unsigned *ids,
                      const size_t *lens, const hsb_lit_ext *ext, unsigned elements,
                      const char *const *regexes, const unsigned *regex_flags, const unsigned *regex_ids, unsigned n_regexes,
                      unsigned mode, const hsb_lit_options *options, const hsb_lit_build_ext *build_ext,
                      unsigned regex_engine, unsigned dfa_states, hs_database_t **db, char **error) {
    if (error) *error = nullptr;
    if (!db) return HSB_LIT_COMPILER_ERROR;
    *db = nullptr;
    std::string err;
    auto fail = [&](const std::string &m) { if (error) *error = dup_error(m); return HSB_LIT_COMPILER_ERROR; };
    if (!expressions || !lens || !elements) return fail("invalid parameter: no expressions");
    if (mode != hsb::MODE_BLOCK && mode != hsb::MODE_VECTORED && mode != hsb::MODE_STREAM) return fail("unknown mode");
    hsb_lit_options opt{};
    if (options) opt = *options;
    BuildExt bx;
    if (build_ext) {
        bx.included_jump = build_ext->included_jump != 0;
        bx.avx512 = build_ext->target_avx512 != 0;
        bx.groups_off = build_ext->groups_off;
        bx.max_bi_anchored_width = build_ext->max_bi_anchored_width;
    }
    bx.full_rose = true;  // literal matches go through the full roseRunProgram (src/rose/match.c:479-523 -> program_runtime.c:2241)
    bx.regex_engine = regex_engine;
    bx.dfa_states = dfa_states;
    for (unsigned i = 0; i < n_regexes; i++) {
        if (!regexes || !regexes[i]) return fail("null regular expression");
        bx.regexes.push_back({regexes[i], regex_flags ? regex_flags[i] : 0, regex_ids ? regex_ids[i] : 0});
    }
    if (build_ext && build_ext->reserved_pure_literal) bx.full_rose = !bx.regexes.empty();
    bx.vectored = mode == hsb::MODE_VECTORED;
    bx.streaming = mode == hsb::MODE_STREAM;
    if ((bx.vectored || bx.streaming) && bx.full_rose) return fail("HS_MODE_VECTORED / HS_MODE_STREAM are written for pure-literal databases only");
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
        if (ext) {
            const hsb_lit_ext &e = ext[i];
            fl.group = e.group;
            fl.check_groups = (e.flags & HSB_EXT_CHECK_GROUPS) != 0;
            fl.want_squash = (e.flags & HSB_EXT_SQUASH_GROUP) != 0;
            fl.quiet = (e.flags & HSB_EXT_QUIET) != 0;
            for (unsigned k = 0; k < e.look_count; k++) {
                Look l{};
                l.offset = e.look[k].offset;
                l.is_class = e.look[k].cls != nullptr;
                l.andm = e.look[k].and_mask;
                l.cmpm = e.look[k].cmp_mask;
                l.neg = e.look[k].negate;
                if (l.is_class) std::memcpy(l.cls, e.look[k].cls, 32);
                fl.look.push_back(l);
            }
        }
        fulls.push_back(fl);
    }
    std::vector<u8> image;
    int rv = build(fulls, opt, bx, image, err);
    if (rv != HSB_LIT_SUCCESS) return fail(err);
    void *mem = nullptr;
    if (posix_memalign(&mem, 64, image.size())) return fail("out of memory");
    std::memcpy(mem, image.data(), image.size());
    *db = (hs_database_t *)mem;
    return HSB_LIT_SUCCESS;
}

int hsb_compile_lit_multi_ext(const char *const *expressions, const unsigned *flags, const unsigned *ids,
                              const size_t *lens, const hsb_lit_ext *ext, unsigned elements, unsigned mode,
                              const hsb_lit_options *options, const hsb_lit_build_ext *build_ext,
                              hs_database_t **db, char **error) {
    hsb_lit_build_ext bx{};
    if (build_ext) bx = *build_ext;
    bx.reserved_pure_literal = 1;
    return hsb_compile_mixed(expressions, flags, ids, lens, ext, elements, nullptr, nullptr, nullptr, 0, mode, options, &bx,
                             0, 0, db, error);
}

int hsb_compile_lit_multi(const char *const *expressions, const unsigned *flags, const unsigned *ids,
                          const size_t *lens, unsigned elements, unsigned mode,
                          const hsb_lit_options *options, hs_database_t **db, char **error) {
    return hsb_compile_lit_multi_ext(expressions, flags, ids, lens, nullptr, elements, mode, options, nullptr, db, error);
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
