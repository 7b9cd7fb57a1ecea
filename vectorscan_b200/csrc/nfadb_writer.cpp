// nfadb_writer.cpp -- host-side writer of SINGLE-OUTFIX block-mode databases whose one engine is a LimEx NFA,
// in the Vectorscan 5.4.11 bytecode format (part of libhs_litdb.so).
//
// Why this exists: the reference's pattern compiler (hs_compile_multi) cannot be built in this image
// (SURVEY.md section 0.5), so there is no source of databases that contain an NFA.  This file compiles a small
// regular-expression dialect into ONE Glushkov automaton and lays it out as the reference does for a pattern
// set that becomes a single outfix: hs_database header, RoseEngine with runtimeImpl = SINGLE_OUTFIX
// (src/runtime.c:246-279 runs it), one NfaInfo, one report program per pattern and a LimExNFA{32..512}
// (src/nfa/limex_internal.h) behind a struct NFA header.  The unmodified reference runtime accepts the blob
// and scans it with hs_scan; the GPU runtime steps the very same bytes.
//
// Dialect (bytes, no UTF-8): literals, escapes \n \r \t \f \v \a \e \0 \xHH \d \D \w \W \s \S and escaped
// punctuation, '.', classes [a-z0-9_] [^...] (escapes allowed inside), groups ( ) and (?: ), alternation |,
// quantifiers * + ? {m} {m,} {m,n} (greedy/lazy makes no difference to the set of end offsets), a leading ^.
// Flags: HS_FLAG_CASELESS, HS_FLAG_DOTALL.  Semantics = Hyperscan's: every end offset of every non-empty match
// is reported once per pattern.  Patterns that match the empty string are refused (they would need the
// boundary programs), as is anything that does not fit 510 positions.
//
// Encoding choices (any LimEx that follows the format is valid input for the runtime): state 0 is the
// unanchored start (self loop), the next state the anchored start, then the Glushkov positions of each pattern
// in order; the <= 8 most frequent forward distances (0..16) between a state and its successors become the
// shift masks, every other edge and every reporting state becomes an exception; no acceleration, no bounded
// repeats, no squashing.
//
// Format sources followed (restated, nothing copied): src/nfa/limex_internal.h:33-201, nfa_internal.h:89-128,
// src/nfa/limex_runtime_impl.h:209-367,620-730 (what the runtime reads), src/rose/rose_internal.h:153-164,330-483,
// src/rose/rose_build_bytecode.cpp:3608-3890 (engine header of a single-outfix build).
#include "hsb_layout.h"
#include "hs_litdb.h"

#include <algorithm>
#include <bitset>
#include <cstdlib>
#include <map>
#include <set>
#include <string>
#include <unordered_map>
#include <vector>

namespace {
using namespace hsb;

constexpr u32 MAX_STATES = 512;
using StateSet = std::bitset<MAX_STATES>;

struct CharSet {
    u64 w[4] = {0, 0, 0, 0};
    void set(u32 c) { w[c >> 6] |= 1ull << (c & 63); }
    bool test(u32 c) const { return (w[c >> 6] >> (c & 63)) & 1; }
    void set_range(u32 a, u32 b) { for (u32 c = a; c <= b; c++) set(c); }
    void invert() { for (auto &x : w) x = ~x; }
    void merge(const CharSet &o) { for (int i = 0; i < 4; i++) w[i] |= o.w[i]; }
    bool empty() const { return !(w[0] | w[1] | w[2] | w[3]); }
    void add_caseless() {
        for (u32 c = 'a'; c <= 'z'; c++) {
            if (test(c)) set(c - 32);
            if (test(c - 32)) set(c);
        }
    }
};

struct Buf {
    std::vector<u8> b;
    size_t size() const { return b.size(); }
    void align(size_t a) { b.resize(round_up(b.size(), a), 0); }
    size_t grow(size_t n) { size_t o = b.size(); b.resize(o + n, 0); return o; }
    u8 *at(size_t off) { return b.data() + off; }
};

// ---- regular expression -> syntax tree --------------------------------------------------------------------
enum NodeType { N_EMPTY, N_CHAR, N_CAT, N_ALT, N_STAR, N_PLUS, N_OPT };
struct Node {
    NodeType type = N_EMPTY;
    CharSet cs;
    std::vector<int> kids;
};

struct Parser {
    const std::string &src;
    bool caseless, dotall;
    size_t pos = 0;
    std::vector<Node> nodes;
    std::string err;
    size_t n_char_nodes = 0;

    Parser(const std::string &s, bool cl, bool da) : src(s), caseless(cl), dotall(da) {}
    bool eof() const { return pos >= src.size(); }
    u8 peek() const { return (u8)src[pos]; }
    int add(NodeType t) { nodes.push_back(Node()); nodes.back().type = t; return (int)nodes.size() - 1; }
    bool fail(const std::string &m) { if (err.empty()) err = m + " at offset " + std::to_string(pos); return false; }

    int make_char(CharSet cs) {
        if (caseless) cs.add_caseless();
        int n = add(N_CHAR);
        nodes[n].cs = cs;
        n_char_nodes++;
        return n;
    }
    int clone(int n) {
        Node copy = nodes[n];
        if (copy.type == N_CHAR) n_char_nodes++;
        for (int &k : copy.kids) k = clone(k);
        nodes.push_back(copy);
        return (int)nodes.size() - 1;
    }

    static int hexval(u8 c) {
        if (c >= '0' && c <= '9') return c - '0';
        if (c >= 'a' && c <= 'f') return c - 'a' + 10;
        if (c >= 'A' && c <= 'F') return c - 'A' + 10;
        return -1;
    }
    // escape after the backslash; fills cs (single byte or a class)
    bool parse_escape(CharSet &cs) {
        if (eof()) return fail("trailing backslash");
        const u8 c = (u8)src[pos++];
        switch (c) {
        case 'n': cs.set('\n'); return true;
        case 'r': cs.set('\r'); return true;
        case 't': cs.set('\t'); return true;
        case 'f': cs.set('\f'); return true;
        case 'v': cs.set('\v'); return true;
        case 'a': cs.set(7); return true;
        case 'e': cs.set(27); return true;
        case '0': cs.set(0); return true;
        case 'x': {
            if (pos + 2 > src.size() || hexval((u8)src[pos]) < 0 || hexval((u8)src[pos + 1]) < 0) return fail("bad \\x escape");
            cs.set((u32)(hexval((u8)src[pos]) * 16 + hexval((u8)src[pos + 1])));
            pos += 2;
            return true;
        }
        case 'd': cs.set_range('0', '9'); return true;
        case 'D': cs.set_range('0', '9'); cs.invert(); return true;
        case 'w': cs.set_range('0', '9'); cs.set_range('a', 'z'); cs.set_range('A', 'Z'); cs.set('_'); return true;
        case 'W': cs.set_range('0', '9'); cs.set_range('a', 'z'); cs.set_range('A', 'Z'); cs.set('_'); cs.invert(); return true;
        case 's': cs.set(' '); cs.set_range(9, 13); return true;
        case 'S': cs.set(' '); cs.set_range(9, 13); cs.invert(); return true;
        default:
            if ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '1' && c <= '9'))
                return fail("unsupported escape");
            cs.set(c);
            return true;
        }
    }
    bool parse_class(CharSet &cs) {  // after '['
        bool negate = false;
        if (!eof() && peek() == '^') { negate = true; pos++; }
        bool first = true;
        for (;;) {
            if (eof()) return fail("unterminated class");
            u8 c = (u8)src[pos++];
            if (c == ']' && !first) break;
            first = false;
            CharSet item;
            bool single = true;
            u32 lo = c;
            if (c == '\\') {
                const size_t before = pos;
                if (!parse_escape(item)) return false;
                const u8 e = (u8)src[before];
                single = !(e == 'd' || e == 'D' || e == 'w' || e == 'W' || e == 's' || e == 'S');
                if (single) for (u32 k = 0; k < 256; k++) if (item.test(k)) lo = k;
            } else {
                item.set(c);
            }
            if (single && pos + 1 < src.size() && src[pos] == '-' && src[pos + 1] != ']') {
                pos++;
                u8 h = (u8)src[pos++];
                u32 hi = h;
                if (h == '\\') {
                    CharSet hs;
                    if (!parse_escape(hs)) return false;
                    int cnt = 0;
                    for (u32 k = 0; k < 256; k++) if (hs.test(k)) { hi = k; cnt++; }
                    if (cnt != 1) return fail("bad range end");
                }
                if (hi < lo) return fail("reversed range");
                item = CharSet();
                item.set_range(lo, hi);
            }
            cs.merge(item);
        }
        if (caseless) cs.add_caseless();
        if (negate) cs.invert();
        if (cs.empty()) return fail("empty class");
        return true;
    }

    int parse_alt() {
        std::vector<int> alts;
        int c = parse_cat();
        if (c < 0) return -1;
        alts.push_back(c);
        while (!eof() && peek() == '|') {
            pos++;
            c = parse_cat();
            if (c < 0) return -1;
            alts.push_back(c);
        }
        if (alts.size() == 1) return alts[0];
        int n = add(N_ALT);
        nodes[n].kids = alts;
        return n;
    }
    int parse_cat() {
        std::vector<int> items;
        while (!eof() && peek() != '|' && peek() != ')') {
            int r = parse_repeat();
            if (r < 0) return -1;
            items.push_back(r);
        }
        if (items.empty()) return add(N_EMPTY);
        if (items.size() == 1) return items[0];
        int n = add(N_CAT);
        nodes[n].kids = items;
        return n;
    }
    int wrap(NodeType t, int kid) { int n = add(t); nodes[n].kids = {kid}; return n; }
    int parse_repeat() {
        int a = parse_atom();
        if (a < 0) return -1;
        for (;;) {
            if (eof()) break;
            const u8 c = peek();
            if (c == '*') { pos++; a = wrap(N_STAR, a); }
            else if (c == '+') { pos++; a = wrap(N_PLUS, a); }
            else if (c == '?') { pos++; a = wrap(N_OPT, a); }
            else if (c == '{') {
                size_t p = pos + 1;
                auto number = [&](u32 &v) { size_t s0 = p; v = 0; while (p < src.size() && src[p] >= '0' && src[p] <= '9') { v = v * 10 + (u32)(src[p] - '0'); p++; if (v > 1000) return false; } return p > s0; };
                u32 lo = 0, hi = 0;
                bool open_ended = false;
                if (!number(lo)) { fail("bad repeat"); return -1; }
                if (p < src.size() && src[p] == ',') {
                    p++;
                    if (p < src.size() && src[p] == '}') open_ended = true;
                    else if (!number(hi)) { fail("bad repeat"); return -1; }
                } else {
                    hi = lo;
                }
                if (p >= src.size() || src[p] != '}') { fail("bad repeat"); return -1; }
                pos = p + 1;
                if (!open_ended && hi < lo) { fail("reversed repeat bounds"); return -1; }
                if (!open_ended && hi == 0) { a = add(N_EMPTY); continue; }
                // expand: lo copies, then (hi - lo) nested optionals, or a star for {lo,}
                std::vector<int> seq;
                for (u32 i = 0; i < lo; i++) seq.push_back(i == 0 ? a : clone(a));
                if (open_ended) {
                    seq.push_back(wrap(N_STAR, lo == 0 ? a : clone(a)));
                } else if (hi > lo) {
                    int tail = -1;  // (a (a (a)?)?)?
                    for (u32 i = 0; i < hi - lo; i++) {
                        int copy = (lo == 0 && i == 0) ? a : clone(a);
                        if (tail < 0) tail = wrap(N_OPT, copy);
                        else { int cat = add(N_CAT); nodes[cat].kids = {copy, tail}; tail = wrap(N_OPT, cat); }
                    }
                    seq.push_back(tail);
                }
                if (n_char_nodes > MAX_STATES) { fail("pattern too large"); return -1; }
                if (seq.size() == 1) a = seq[0];
                else { int cat = add(N_CAT); nodes[cat].kids = seq; a = cat; }
            } else break;
            if (!eof() && (peek() == '?' || peek() == '+') && (c == '*' || c == '+' || c == '?' || c == '{')) pos++;  // lazy / possessive marker
        }
        return a;
    }
    int parse_atom() {
        if (eof()) { fail("unexpected end"); return -1; }
        const u8 c = (u8)src[pos++];
        CharSet cs;
        switch (c) {
        case '(': {
            if (pos + 1 < src.size() && src[pos] == '?' && src[pos + 1] == ':') pos += 2;
            else if (!eof() && peek() == '?') { fail("unsupported group syntax"); return -1; }
            int n = parse_alt();
            if (n < 0) return -1;
            if (eof() || peek() != ')') { fail("missing )"); return -1; }
            pos++;
            return n;
        }
        case '[':
            if (!parse_class(cs)) return -1;
            { int n = add(N_CHAR); nodes[n].cs = cs; n_char_nodes++; return n; }
        case '.':
            cs.invert();
            if (!dotall) { CharSet nl; nl.set('\n'); for (int i = 0; i < 4; i++) cs.w[i] &= ~nl.w[i]; }
            { int n = add(N_CHAR); nodes[n].cs = cs; n_char_nodes++; return n; }
        case '\\':
            if (!parse_escape(cs)) return -1;
            return make_char(cs);
        case '*': case '+': case '?': case '{':
            fail("nothing to repeat");
            return -1;
        case '^': case '$':
            fail("anchors are only supported as a leading ^");
            return -1;
        default:
            cs.set(c);
            return make_char(cs);
        }
    }
};

// ---- Glushkov construction ----------------------------------------------------------------------------------
struct Glushkov {
    std::vector<CharSet> cls;        // per position
    std::vector<StateSet> follow;    // per position
    struct Info { bool nullable; StateSet first, last; u32 min_len; };

    Info build(const std::vector<Node> &nodes, int n) {
        const Node &nd = nodes[n];
        Info r{true, {}, {}, 0};
        switch (nd.type) {
        case N_EMPTY:
            return r;
        case N_CHAR: {
            const u32 p = (u32)cls.size();
            cls.push_back(nd.cs);
            follow.push_back(StateSet());
            r.nullable = false;
            r.first.set(p);
            r.last.set(p);
            r.min_len = 1;
            return r;
        }
        case N_CAT: {
            bool have = false;
            for (int k : nd.kids) {
                Info ki = build(nodes, k);
                if (!have) { r = ki; have = true; continue; }
                for (u32 p = 0; p < cls.size(); p++) if (r.last.test(p)) follow[p] |= ki.first;
                if (r.nullable) r.first |= ki.first;
                if (ki.nullable) r.last |= ki.last; else r.last = ki.last;
                r.nullable = r.nullable && ki.nullable;
                r.min_len += ki.min_len;
            }
            return r;
        }
        case N_ALT: {
            bool have = false;
            for (int k : nd.kids) {
                Info ki = build(nodes, k);
                if (!have) { r = ki; have = true; continue; }
                r.nullable = r.nullable || ki.nullable;
                r.first |= ki.first;
                r.last |= ki.last;
                r.min_len = std::min(r.min_len, ki.min_len);
            }
            return r;
        }
        case N_STAR: case N_PLUS: case N_OPT: {
            Info ki = build(nodes, nd.kids[0]);
            r = ki;
            if (nd.type != N_OPT)
                for (u32 p = 0; p < cls.size(); p++) if (ki.last.test(p)) follow[p] |= ki.first;
            if (nd.type != N_PLUS) { r.nullable = true; r.min_len = 0; }
            return r;
        }
        }
        return r;
    }
};

}  // namespace

#include "outfix_engines.h"
namespace hsb_nfadb {
int build(const std::vector<Pattern> &pats, unsigned engine, std::vector<hsb::u8> &image, std::string &err, unsigned mode = hsb::MODE_BLOCK);
int build_rose(const std::vector<Pattern> &pats, unsigned engine, hsb::u32 dfa_states, std::vector<hsb::u8> &image, std::string &err);
}

namespace {

// little-endian bit vector of `bytes` bytes
void put_state(u8 *dst, const StateSet &s, u32 bytes) {
    std::memset(dst, 0, bytes);
    for (u32 i = 0; i < bytes * 8 && i < MAX_STATES; i++)
        if (s.test(i)) dst[i >> 3] |= (u8)(1u << (i & 7));
}

struct Automaton {
    u32 n_states = 0;
    std::vector<CharSet> reach_of;     // bytes on which a state may be entered
    std::vector<StateSet> succ;        // successors of each state
    std::vector<u32> report_of;        // pattern index + 1 a state reports, 0 = none
    std::vector<u32> report_eod_of;    // same, for patterns ending in \z: reported only at the end of the data
    bool any_eod = false;
    StateSet init, init_ds;
    u32 min_width = ~0u;
};


// ---- determinisation: Automaton -> DFA (the McClellan engine, src/nfa/mcclellan_internal.h) -----------------
// Symbols are the byte classes of the reach table (bytes no state tells apart share a column).  State 0 is the
// dead state (empty subset), as in the reference (mcclellancompile.cpp:558 "dead is always 0").
struct Dfa {
    u32 n_sym = 0;
    std::vector<std::vector<u32>> next;        // [state][symbol]
    std::vector<std::vector<u32>> reports;     // [state] sorted pattern indices accepted on entering the state
    std::vector<std::vector<u32>> reports_eod; // [state] ... accepted only if the data ends in this state
    u32 start_anchored = 0, start_floating = 0;
};

struct StateSetHash {
    size_t operator()(const StateSet &s) const { return std::hash<StateSet>()(s); }
};

bool determinise(const Automaton &am, const std::vector<StateSet> &reach_rows, Dfa &d, u32 max_states) {
    d.n_sym = (u32)reach_rows.size();
    std::unordered_map<StateSet, u32, StateSetHash> id_of;
    std::vector<StateSet> subset;
    auto intern = [&](const StateSet &ss) -> u32 {
        auto it = id_of.find(ss);
        if (it != id_of.end()) return it->second;
        const u32 id = (u32)subset.size();
        id_of.emplace(ss, id);
        subset.push_back(ss);
        return id;
    };
    intern(StateSet());  // dead
    d.start_anchored = intern(am.init);
    d.start_floating = intern(am.init_ds);
    for (u32 cur = 0; cur < subset.size(); cur++) {
        if (subset.size() > max_states) return false;
        const StateSet ss = subset[cur];
        StateSet all_succ;
        for (u32 s = 0; s < am.n_states; s++) if (ss.test(s)) all_succ |= am.succ[s];
        std::vector<u32> row(d.n_sym, 0);
        for (u32 k = 0; k < d.n_sym; k++) row[k] = intern(all_succ & reach_rows[k]);
        d.next.push_back(std::move(row));
    }
    if (subset.size() > max_states) return false;
    d.reports.assign(subset.size(), {});
    d.reports_eod.assign(subset.size(), {});
    for (u32 i = 0; i < subset.size(); i++) {
        std::set<u32> r;
        for (u32 s = 0; s < am.n_states; s++) if (subset[i].test(s) && am.report_of[s]) r.insert(am.report_of[s] - 1);
        d.reports[i].assign(r.begin(), r.end());
        std::set<u32> re;
        for (u32 s = 0; s < am.n_states; s++) if (subset[i].test(s) && am.report_eod_of[s]) re.insert(am.report_eod_of[s] - 1);
        d.reports_eod[i].assign(re.begin(), re.end());
    }
    return true;
}

// Moore partition refinement; keeps state 0 dead and renumbers in discovery order from the start states.
void minimise(Dfa &d) {
    const u32 n = (u32)d.next.size();
    std::vector<u32> cls(n);
    {
        std::map<std::pair<std::vector<u32>, std::vector<u32>>, u32> by_rep;
        for (u32 i = 0; i < n; i++) {
            const auto key = std::make_pair(d.reports[i], d.reports_eod[i]);
            auto it = by_rep.find(key);
            if (it == by_rep.end()) it = by_rep.emplace(key, (u32)by_rep.size()).first;
            cls[i] = it->second;
        }
    }
    for (;;) {
        std::map<std::vector<u32>, u32> sig_id;
        std::vector<u32> ncls(n);
        std::vector<u32> sig(d.n_sym + 1);
        for (u32 i = 0; i < n; i++) {
            sig[0] = cls[i];
            for (u32 k = 0; k < d.n_sym; k++) sig[k + 1] = cls[d.next[i][k]];
            auto it = sig_id.find(sig);
            if (it == sig_id.end()) it = sig_id.emplace(sig, (u32)sig_id.size()).first;
            ncls[i] = it->second;
        }
        u32 before = 0, after = (u32)sig_id.size();
        { std::set<u32> u(cls.begin(), cls.end()); before = (u32)u.size(); }
        cls.swap(ncls);
        if (after == before) break;
    }
    // renumber: dead class first, then breadth-first from the start states
    std::vector<u32> rep_of_class, new_id;
    {
        u32 ncls = 0;
        for (u32 c : cls) ncls = std::max(ncls, c + 1);
        rep_of_class.assign(ncls, ~0u);
        for (u32 i = 0; i < n; i++) if (rep_of_class[cls[i]] == ~0u) rep_of_class[cls[i]] = i;
        new_id.assign(ncls, ~0u);
    }
    std::vector<u32> order;
    auto visit = [&](u32 c) { if (new_id[c] == ~0u) { new_id[c] = (u32)order.size(); order.push_back(c); } };
    visit(cls[0]);
    visit(cls[d.start_anchored]);
    visit(cls[d.start_floating]);
    for (u32 q = 0; q < order.size(); q++) {
        const u32 r = rep_of_class[order[q]];
        for (u32 k = 0; k < d.n_sym; k++) visit(cls[d.next[r][k]]);
    }
    Dfa m;
    m.n_sym = d.n_sym;
    m.next.resize(order.size());
    m.reports.resize(order.size());
    m.reports_eod.resize(order.size());
    for (u32 q = 0; q < order.size(); q++) {
        const u32 r = rep_of_class[order[q]];
        m.next[q].resize(d.n_sym);
        for (u32 k = 0; k < d.n_sym; k++) m.next[q][k] = new_id[cls[d.next[r][k]]];
        m.reports[q] = d.reports[r];
        m.reports_eod[q] = d.reports_eod[r];
    }
    m.start_anchored = new_id[cls[d.start_anchored]];
    m.start_floating = new_id[cls[d.start_floating]];
    d = std::move(m);
}


// Lay the DFA out as the reference's McClellan engine (struct NFA + struct mcclellan + successor table +
// mstate_aux[] + report lists [+ Sherman states]), following mcclellancompile.cpp:611-880 (16-bit) and :924-1000
// (8-bit).  ReportIDs are Rose report-program offsets, as for every outfix engine (roseReportAdaptor).
// `sherman`: compress states that differ from the start state or from their discovery parent in at most
// MAX_SHERMAN_LIST_LEN symbols into 32-byte Sherman states (16-bit engines only).
// `wide_states` (16-bit engines): chains of at least 8 non-accepting states, each entered only from its predecessor and
// all leaving the chain the same way (one failure row), are stored as ONE wide head state -- the chain's symbols, the
// successor after the last one, the shared failure row (mcclellancompile.cpp:773-838, struct layout
// mcclellan_internal.h:53-61, runtime doWide16 mcclellan_common_impl.h:86-188) -- the reference's compression of
// literal-like runs.  Chain states have no table row and no mstate_aux.
int emit_mcclellan(const Dfa &d, const u8 *reach_map, const std::vector<u32> &program_of, bool sherman,
                   u32 min_width, Buf &eng, u32 &state_bytes, std::string &err, bool wide_states = false) {
    namespace mc = mcclellan;
    const u32 n = (u32)d.next.size();
    const u32 alpha = d.n_sym + 1;  // + TOP, the last symbol (never produced by remap)
    u32 as = 0;
    while ((1u << as) < alpha) as++;
    const bool wide16 = n > 256;
    if (n - 1 > mc::STATE_MASK) { err = "DFA too large for a McClellan engine (" + std::to_string(n) + " states)"; return HSB_LIT_COMPILER_ERROR; }
    state_bytes = wide16 ? 2 : 1;
    // ---- implementation ids ----
    std::vector<u32> impl(n, 0);
    std::vector<u32> daddy(n, 0);
    std::vector<u8> is_sherman(n, 0);
    u32 accel_limit_8 = 0, accept_limit_8 = 0, sherman_limit = 0, n_real = n, wide_limit = 0;
    // ---- wide chains ----
    std::vector<std::vector<u32>> chain_states, chain_syms;
    std::vector<u32> chain_target;    // successor after the chain's last symbol
    std::vector<std::vector<u32>> chain_fail;   // the shared failure row, one successor per symbol
    std::vector<u8> in_chain(n, 0);   // 1 head, 2 inner
    if (wide16 && wide_states) {
        // inner candidates: non-accepting, not a start state, entered by exactly one (state, symbol) edge and not named
        // as the successor of an earlier chain (`pinned`: such a state must keep its table row)
        std::vector<u32> in_edges(n, 0), pred(n, 0);
        std::vector<u8> pinned(n, 0);
        for (u32 q = 0; q < n; q++)
            for (u32 k = 0; k < d.n_sym; k++) { const u32 t = d.next[q][k]; in_edges[t]++; pred[t] = q; }
        auto plain = [&](u32 q) { return q && d.reports[q].empty() && d.reports_eod[q].empty() && q != d.start_anchored && q != d.start_floating; };
        for (u32 h = 1; h < n; h++) {
            if (!plain(h) || in_chain[h]) continue;
            std::vector<u32> st{h}, sy;
            std::vector<int> F(d.n_sym, -1);   // successor on a symbol that is not the state's own chain symbol
            u32 cur = h;
            while (st.size() < 4096) {
                int own = -1;
                for (u32 k = 0; k < d.n_sym && own < 0; k++) {
                    const u32 t = d.next[cur][k];
                    if (plain(t) && in_edges[t] == 1 && pred[t] == cur && !in_chain[t] && !pinned[t] && t != h) own = (int)k;
                }
                if (own < 0) break;
                bool consistent = true;
                for (u32 k = 0; k < d.n_sym && consistent; k++)
                    if (k != (u32)own && F[k] >= 0 && (u32)F[k] != d.next[cur][k]) consistent = false;
                if (!consistent) break;
                for (u32 k = 0; k < d.n_sym; k++) if (k != (u32)own) F[k] = (int)d.next[cur][k];
                sy.push_back((u32)own);
                cur = d.next[cur][own];
                st.push_back(cur);
            }
            const u32 w = (u32)sy.size();   // st holds w chain states and, last, the state the chain leads to
            if (w < 8) continue;
            const u32 target = st[w];
            st.resize(w);
            pinned[target] = 1;
            in_chain[st[0]] = 1;
            for (u32 j = 1; j < w; j++) in_chain[st[j]] = 2;
            std::vector<u32> fail(d.n_sym);
            for (u32 k = 0; k < d.n_sym; k++) fail[k] = F[k] >= 0 ? (u32)F[k] : target;   // mcclellancompile.cpp:815-829
            chain_states.push_back(st);
            chain_syms.push_back(sy);
            chain_target.push_back(target);
            chain_fail.push_back(fail);
        }
    }
    if (!wide16) {
        u32 j = 1;
        for (u32 i = 1; i < n; i++) if (d.reports[i].empty()) impl[i] = j++;
        accel_limit_8 = accept_limit_8 = j;  // no accel states
        for (u32 i = 1; i < n; i++) if (!d.reports[i].empty()) impl[i] = j++;
    } else {
        if (sherman) {
            std::vector<u32> parent(n, 0);
            std::vector<u8> seen(n, 0);
            seen[0] = 1;
            for (u32 st : {d.start_anchored, d.start_floating}) seen[st] = 1;
            for (u32 q = 0; q < n; q++)  // ids are already in discovery order (minimise)
                for (u32 k = 0; k < d.n_sym; k++) { const u32 t = d.next[q][k]; if (!seen[t]) { seen[t] = 1; parent[t] = q; } }
            for (u32 i = 1; i < n; i++) {
                if (i == d.start_anchored || i == d.start_floating || in_chain[i]) continue;
                for (u32 cand : {d.start_floating ? d.start_floating : d.start_anchored, parent[i]}) {
                    if (!cand || cand == i || is_sherman[cand] || in_chain[cand]) continue;   // a daddy needs a table row
                    u32 diff = 0;
                    for (u32 k = 0; k < d.n_sym; k++) diff += d.next[i][k] != d.next[cand][k];
                    if (diff <= mc::MAX_SHERMAN_LIST_LEN) { is_sherman[i] = 1; daddy[i] = cand; break; }
                }
            }
            // a daddy must be a normal state (doSherman16 reads its row): ids are in discovery order, so a parent is
            // decided before its children; this pass only guards that invariant
            for (u32 i = 1; i < n; i++) if (is_sherman[i] && is_sherman[daddy[i]]) is_sherman[i] = 0;
        }
        u32 j = 1;   // allocateFSN16: normal, Sherman, wide heads, the other chain states
        for (u32 i = 1; i < n; i++) if (!is_sherman[i] && !in_chain[i]) impl[i] = j++;
        sherman_limit = n_real = j;
        for (u32 i = 1; i < n; i++) if (is_sherman[i]) impl[i] = j++;
        wide_limit = j;
        for (auto &c : chain_states) impl[c[0]] = j++;
        for (auto &c : chain_states) for (size_t q = 1; q < c.size(); q++) impl[c[q]] = j++;
    }
    const bool has_wide = !chain_states.empty();
    const u32 n_aux = has_wide ? wide_limit : n;   // wide states have no mstate_aux (mcclellancompile.cpp:641)
    // ---- report lists ----
    std::map<std::vector<u32>, u32> list_id;  // sorted program offsets -> index
    std::vector<std::vector<u32>> lists;
    std::vector<u32> list_of(n, ~0u);
    std::set<u32> all_reports;
    for (u32 i = 0; i < n; i++) {
        if (d.reports[i].empty()) continue;
        std::set<u32> ps;
        for (u32 pi : d.reports[i]) ps.insert(program_of[pi]);
        std::vector<u32> l(ps.begin(), ps.end());
        all_reports.insert(l.begin(), l.end());
        auto it = list_id.find(l);
        if (it == list_id.end()) { it = list_id.emplace(l, (u32)lists.size()).first; lists.push_back(l); }
        list_of[i] = it->second;
    }
    std::vector<u32> list_eod_of(n, ~0u);
    bool any_eod = false;
    for (u32 i = 0; i < n; i++) {
        if (d.reports_eod[i].empty()) continue;
        any_eod = true;
        std::set<u32> ps;
        for (u32 pi : d.reports_eod[i]) ps.insert(program_of[pi]);
        std::vector<u32> l(ps.begin(), ps.end());
        all_reports.insert(l.begin(), l.end());
        auto it = list_id.find(l);
        if (it == list_id.end()) { it = list_id.emplace(l, (u32)lists.size()).first; lists.push_back(l); }
        list_eod_of[i] = it->second;
    }
    const bool single = all_reports.size() == 1;
    // ---- sizes (mcclellancompile.cpp:637-652 / :934-940) ----
    const size_t ent = wide16 ? 2 : 1;
    const size_t tran_size = ((size_t)1 << as) * ent * n_real;
    const size_t aux_size = mc::aux::SIZE * (size_t)n_aux;
    const size_t aux_offset = round_up(nfa::SIZE + mc::SIZE + tran_size, 16);
    size_t rl_size = 0;
    for (auto &l : lists) rl_size += 4 + 4 * l.size();
    const size_t accel_offset = round_up(aux_offset + aux_size + rl_size, 32);
    const size_t sherman_offset = round_up(accel_offset, 16);
    const size_t sherman_size = wide16 ? (size_t)mc::SHERMAN_FIXED_SIZE * ((has_wide ? wide_limit : n) - n_real) : 0;
    const size_t wide_offset = round_up(sherman_offset + sherman_size, 16);
    size_t wide_size = 0;   // calcWideRegionSize, mcclellancompile.cpp:522-539
    if (has_wide) {
        wide_size = 4 + 4 * chain_states.size();
        for (auto &c : chain_syms) wide_size += round_up(c.size(), 2) + 2 * ((size_t)d.n_sym + 1) + 2;
        wide_size = round_up(wide_size, 16);
    }
    const size_t total = round_up(wide_offset + wide_size, 16);
    eng.grow(total);
    eng.align(CACHELINE);
    u8 *nh = eng.at(0), *m = eng.at(nfa::SIZE);
    wr32(nh, nfa::flags, any_eod ? nfa::ACCEPTS_EOD : 0);  // mcclellancompile.cpp:1513-1515
    wr32(nh, nfa::length, (u32)eng.size());
    wr8(nh, nfa::type, wide16 ? nfa::MCCLELLAN_16 : nfa::MCCLELLAN_8);
    wr32(nh, nfa::queueIndex, 0);
    wr32(nh, nfa::nPositions, n);
    if (has_wide) state_bytes += 2;   // the offset inside a wide state is kept beside the state (mcclellancompile.cpp:345-349)
    wr32(nh, nfa::scratchStateSize, state_bytes);
    wr32(nh, nfa::streamStateSize, state_bytes);
    wr32(nh, nfa::maxWidth, 0);
    wr32(nh, nfa::minWidth, min_width);
    wr16(m, mc::state_count, (u16)n);
    wr32(m, mc::length, (u32)eng.size());
    wr16(m, mc::start_anchored, (u16)impl[d.start_anchored]);
    wr16(m, mc::start_floating, (u16)impl[d.start_floating]);
    wr32(m, mc::aux_offset, (u32)aux_offset);
    wr32(m, mc::sherman_offset, wide16 ? (u32)sherman_offset : 0);
    wr32(m, mc::sherman_end, wide16 ? (u32)total : 0);
    wr16(m, mc::accel_limit_8, (u16)accel_limit_8);
    wr16(m, mc::accept_limit_8, (u16)accept_limit_8);
    wr16(m, mc::sherman_limit, (u16)sherman_limit);
    wr16(m, mc::wide_limit, has_wide ? (u16)wide_limit : wide16 ? (u16)n : 0);
    wr32(m, mc::wide_offset, has_wide ? (u32)wide_offset : 0);
    wr8(m, mc::alphaShift, (u8)as);
    wr8(m, mc::flags, single ? mc::FLAG_SINGLE : 0);
    wr8(m, mc::has_accel, 0);
    wr8(m, mc::has_wide, has_wide ? 1 : 0);
    std::memcpy(m + mc::remap, reach_map, 256);
    wr32(m, mc::arb_report, all_reports.empty() ? 0 : *all_reports.begin());
    wr32(m, mc::accel_offset, (u32)(accel_offset - nfa::SIZE));
    // report lists behind the aux array
    std::vector<u32> list_off(lists.size());
    {
        size_t o = aux_offset + aux_size;
        for (size_t i = 0; i < lists.size(); i++) {
            list_off[i] = (u32)o;
            wr32(eng.at(o), 0, (u32)lists[i].size());
            for (size_t k = 0; k < lists[i].size(); k++) wr32(eng.at(o), 4 + 4 * k, lists[i][k]);
            o += 4 + 4 * lists[i].size();
        }
    }
    // aux + successor table
    u8 *succ = m + mc::SIZE;
    auto flagged = [&](u32 t) -> u32 {  // markEdges, mcclellancompile.cpp:188-284: wide states carry no flags
        u32 v = impl[t];
        if (wide16 && !in_chain[t] && !d.reports[t].empty()) v |= mc::ACCEPT_FLAG;
        return v;
    };
    if (has_wide) {
        u8 *wb = eng.at(wide_offset);
        wb[0] = 2;  // WIDE_STATE
        wr16(wb, 2, (u16)chain_states.size());
        size_t o = 4 + 4 * chain_states.size();
        for (size_t c = 0; c < chain_states.size(); c++) {   // heads were numbered in this order
            const u32 w = (u32)chain_syms[c].size();
            wr32(wb, 4 + 4 * c, (u32)o);
            wr16(wb, o, (u16)w);
            for (u32 q = 0; q < w; q++) wb[o + 2 + q] = (u8)chain_syms[c][q];
            size_t t = o + 2 + round_up(w, 2);
            wr16(wb, t, (u16)flagged(chain_target[c]));
            for (u32 k = 0; k < d.n_sym; k++) wr16(wb, t + 2 + 2 * (size_t)k, (u16)flagged(chain_fail[c][k]));
            o = t + 2 * ((size_t)d.n_sym + 1);
        }
    }
    for (u32 i = 0; i < n; i++) {
        if (in_chain[i]) continue;   // no row, no aux
        const u32 id = impl[i];
        u8 *a = eng.at(aux_offset + mc::aux::SIZE * (size_t)id);
        wr32(a, mc::aux::accept, list_of[i] == ~0u ? 0 : list_off[list_of[i]]);
        wr32(a, mc::aux::accept_eod, list_eod_of[i] == ~0u ? 0 : list_off[list_eod_of[i]]);
        wr16(a, mc::aux::top, (u16)impl[i ? d.start_floating : d.start_floating]);
        wr32(a, mc::aux::accel_offset, 0);
        if (wide16 && is_sherman[i]) {
            u8 *sh = eng.at(sherman_offset + (size_t)mc::SHERMAN_FIXED_SIZE * (id - sherman_limit));
            std::vector<u32> ks;
            for (u32 k = 0; k < d.n_sym; k++) if (d.next[i][k] != d.next[daddy[i]][k]) ks.push_back(k);
            sh[0] = (u8)mc::SHERMAN_STATE;
            sh[mc::SHERMAN_LEN_OFFSET] = (u8)ks.size();
            wr16(sh, mc::SHERMAN_DADDY_OFFSET, (u16)impl[daddy[i]]);
            for (size_t q = 0; q < ks.size(); q++) {
                sh[mc::SHERMAN_CHARS_OFFSET + q] = (u8)ks[q];
                wr16(sh, mc::SHERMAN_CHARS_OFFSET + ks.size() + 2 * q, (u16)flagged(d.next[i][ks[q]]));
            }
            continue;
        }
        for (u32 k = 0; k < d.n_sym; k++) {
            const size_t at = ((size_t)id << as) + k;
            if (wide16) wr16(succ, 2 * at, (u16)flagged(d.next[i][k]));
            else succ[at] = (u8)impl[d.next[i][k]];
        }
    }
    return HSB_LIT_SUCCESS;
}


// Lay the DFA out as the reference's McSheng hybrid (struct NFA + struct mcsheng + successor table of the states outside
// the Sheng region + mstate_aux[] + report lists), the engine the reference compiler tries for every DFA too large for a
// pure Sheng (getDfa, src/rose/rose_build_bytecode.cpp: shengCompile, then mcshengCompile, then mcclellanCompile).
// Format: mcsheng_compile.cpp:155-204 (masks), :372-430 / :1266-1307 (ids), :1007-1084 / :1310-1360 (sizes).
// The region: breadth first from the start state over non-accepting states ("cannot raise callbacks from sheng
// region", mcsheng_compile.cpp:582-586), at most 16; in a 16-bit engine every state entered straight from the region
// must have an id that fits the mask byte, so those states are numbered first and the region shrinks until they fit.
// No Sherman states.  Fails (caller falls back to McClellan) when the start state itself accepts.
// `wide64`: struct mcsheng64 (NFA types 23 / 24, AVX512VBMI targets): a region of up to 64 states, 64-byte masks.
int emit_mcsheng(const Dfa &d, const u8 *reach_map, const std::vector<u32> &program_of, u32 min_width, Buf &eng,
                 u32 &state_bytes, std::string &err, bool wide64 = false) {
    namespace mc = mcclellan;
    namespace ms = mcsheng;
    const u32 n = (u32)d.next.size();
    const u32 alpha = d.n_sym + 1;  // + TOP
    u32 as = 0;
    while ((1u << as) < alpha) as++;
    const bool wide16 = n > 256;
    if (n - 1 > mc::STATE_MASK) { err = "DFA too large for a McSheng engine (" + std::to_string(n) + " states)"; return HSB_LIT_COMPILER_ERROR; }
    state_bytes = wide16 ? 2 : 1;
    const u32 start = d.start_anchored ? d.start_anchored : d.start_floating;
    std::vector<u32> region;
    std::vector<u8> in_region(n, 0), is_succ(n, 0);
    const size_t HDR = wide64 ? ms::SIZE64 : ms::SIZE, MASKS = wide64 ? ms::sheng_succ_masks64 : ms::sheng_masks, STRIDE = wide64 ? 64 : 16;
    for (u32 cap = wide64 ? ms::MAX_SHENG64_STATES : ms::MAX_SHENG_STATES; cap >= 1; cap--) {
        region.clear();
        std::fill(in_region.begin(), in_region.end(), 0);
        std::fill(is_succ.begin(), is_succ.end(), 0);
        std::vector<u32> queue{start};
        std::vector<u8> seen(n, 0);
        seen[0] = 1;
        seen[start] = 1;
        for (size_t h = 0; h < queue.size() && region.size() < cap; h++) {
            const u32 q = queue[h];
            if (!d.reports[q].empty()) continue;
            region.push_back(q);
            in_region[q] = 1;
            for (u32 k = 0; k < d.n_sym; k++) { const u32 t = d.next[q][k]; if (!seen[t]) { seen[t] = 1; queue.push_back(t); } }
        }
        u32 n_succ = 0;
        for (u32 q : region)
            for (u32 k = 0; k < d.n_sym; k++) { const u32 t = d.next[q][k]; if (t && !in_region[t] && !is_succ[t]) { is_succ[t] = 1; n_succ++; } }
        if (!wide16 || region.size() + 1 + n_succ <= 256) break;
    }
    if (region.empty()) { err = "no Sheng region: the start state accepts"; return HSB_LIT_COMPILER_ERROR; }
    const u32 sheng_end = (u32)region.size() + 1;
    // ---- implementation ids: dead 0, region 1 .., then (16-bit) region successors, the rest; (8-bit) non-accepting, accepting ----
    std::vector<u32> impl(n, 0);
    for (u32 i = 0; i < region.size(); i++) impl[region[i]] = i + 1;
    u32 j = sheng_end, accel_limit_8 = 0, accept_limit_8 = 0;
    if (wide16) {
        for (u32 i = 1; i < n; i++) if (!in_region[i] && is_succ[i]) impl[i] = j++;
        for (u32 i = 1; i < n; i++) if (!in_region[i] && !is_succ[i]) impl[i] = j++;
    } else {
        for (u32 i = 1; i < n; i++) if (!in_region[i] && d.reports[i].empty()) impl[i] = j++;
        accel_limit_8 = accept_limit_8 = j;
        for (u32 i = 1; i < n; i++) if (!in_region[i] && !d.reports[i].empty()) impl[i] = j++;
    }
    // ---- report lists ----
    std::map<std::vector<u32>, u32> list_id;
    std::vector<std::vector<u32>> lists;
    std::vector<u32> list_of(n, ~0u), list_eod_of(n, ~0u);
    std::set<u32> all_reports;
    bool any_eod = false;
    auto intern = [&](const std::vector<u32> &pis) {
        std::set<u32> ps;
        for (u32 pi : pis) ps.insert(program_of[pi]);
        std::vector<u32> l(ps.begin(), ps.end());
        all_reports.insert(l.begin(), l.end());
        auto it = list_id.find(l);
        if (it == list_id.end()) { it = list_id.emplace(l, (u32)lists.size()).first; lists.push_back(l); }
        return it->second;
    };
    for (u32 i = 0; i < n; i++) {
        if (!d.reports[i].empty()) list_of[i] = intern(d.reports[i]);
        if (!d.reports_eod[i].empty()) { list_eod_of[i] = intern(d.reports_eod[i]); any_eod = true; }
    }
    const bool single = all_reports.size() == 1;
    // ---- sizes ----
    const size_t ent = wide16 ? 2 : 1;
    const u32 n_normal = n - sheng_end;
    const size_t tran_size = ((size_t)1 << as) * ent * n_normal;
    const size_t aux_size = mc::aux::SIZE * (size_t)n;
    const size_t aux_offset = round_up(nfa::SIZE + HDR + tran_size, 16);
    size_t rl_size = 0;
    for (auto &l : lists) rl_size += 4 + 4 * l.size();
    const size_t accel_offset = round_up(aux_offset + aux_size + rl_size, 32);
    const size_t sherman_offset = round_up(accel_offset, 16);
    const size_t total = sherman_offset;
    eng.grow(total);
    eng.align(CACHELINE);
    u8 *nh = eng.at(0), *m = eng.at(nfa::SIZE);
    wr32(nh, nfa::flags, any_eod ? nfa::ACCEPTS_EOD : 0);
    wr32(nh, nfa::length, (u32)eng.size());
    wr8(nh, nfa::type, wide64 ? (wide16 ? nfa::MCSHENG64_16 : nfa::MCSHENG64_8) : (wide16 ? nfa::MCSHENG_16 : nfa::MCSHENG_8));
    wr32(nh, nfa::queueIndex, 0);
    wr32(nh, nfa::nPositions, n);
    wr32(nh, nfa::scratchStateSize, state_bytes);
    wr32(nh, nfa::streamStateSize, state_bytes);
    wr32(nh, nfa::maxWidth, 0);
    wr32(nh, nfa::minWidth, min_width);
    wr16(m, ms::state_count, (u16)n);
    wr32(m, ms::length, (u32)eng.size());
    wr16(m, ms::start_anchored, (u16)impl[d.start_anchored]);
    wr16(m, ms::start_floating, (u16)impl[d.start_floating]);
    wr32(m, ms::aux_offset, (u32)aux_offset);
    wr32(m, ms::sherman_offset, wide16 ? (u32)sherman_offset : 0);
    wr32(m, ms::sherman_end, wide16 ? (u32)total : 0);
    wr16(m, ms::sheng_end, (u16)sheng_end);
    wr16(m, ms::sheng_accel_limit, (u16)(sheng_end - 1));  // no accelerable Sheng state
    wr16(m, ms::accel_limit_8, (u16)accel_limit_8);
    wr16(m, ms::accept_limit_8, (u16)accept_limit_8);
    wr16(m, ms::sherman_limit, wide16 ? (u16)n : 0);
    wr8(m, ms::alphaShift, (u8)as);
    wr8(m, ms::flags, single ? mc::FLAG_SINGLE : 0);
    wr8(m, ms::has_accel, 0);
    std::memcpy(m + ms::remap, reach_map, 256);
    wr32(m, ms::arb_report, all_reports.empty() ? 0 : *all_reports.begin());
    wr32(m, ms::accel_offset, (u32)(accel_offset - nfa::SIZE));
    // shuffle masks, one per byte value (createShuffleMasks)
    for (u32 c = 0; c < 256; c++) {
        u8 *mask = m + MASKS + STRIDE * (size_t)c;
        for (u32 sid = 0; sid < region.size(); sid++) {
            const u32 t = impl[d.next[region[sid]][reach_map[c]]];
            mask[sid] = (u8)(t == 0 ? sheng_end - 1 : t < sheng_end ? t - 1 : t);
        }
    }
    std::vector<u32> list_off(lists.size());
    {
        size_t o = aux_offset + aux_size;
        for (size_t i = 0; i < lists.size(); i++) {
            list_off[i] = (u32)o;
            wr32(eng.at(o), 0, (u32)lists[i].size());
            for (size_t k = 0; k < lists[i].size(); k++) wr32(eng.at(o), 4 + 4 * k, lists[i][k]);
            o += 4 + 4 * lists[i].size();
        }
    }
    u8 *succ = m + HDR;
    for (u32 i = 0; i < n; i++) {
        const u32 id = impl[i];
        u8 *a = eng.at(aux_offset + mc::aux::SIZE * (size_t)id);
        wr32(a, mc::aux::accept, list_of[i] == ~0u ? 0 : list_off[list_of[i]]);
        wr32(a, mc::aux::accept_eod, list_eod_of[i] == ~0u ? 0 : list_off[list_eod_of[i]]);
        wr16(a, mc::aux::top, (u16)impl[d.start_floating]);
        wr32(a, mc::aux::accel_offset, 0);
        if (id < sheng_end) continue;  // dead and region states have no row
        for (u32 k = 0; k < d.n_sym; k++) {
            const size_t at = ((size_t)(id - sheng_end) << as) + k;
            const u32 t = d.next[i][k];
            if (wide16) wr16(succ, 2 * at, (u16)(impl[t] | (d.reports[t].empty() ? 0 : mc::ACCEPT_FLAG)));
            else succ[at] = (u8)impl[t];
        }
    }
    return HSB_LIT_SUCCESS;
}


// Lay a DFA of at most 16 states out as the reference's Sheng engine (struct NFA + struct sheng + sstate_aux[] +
// report lists), following shengcompile.cpp:337-358 (state byte = id | flags), :420-444 (header), :560-578 (masks)
// and :640-698 (region order and sizes).  State ids are the DFA's own (dead = 0).
// `wide`: the AVX512VBMI forms -- Sheng32 (<= 32 states) or Sheng64 (<= 64 states): 64-byte succ_masks, other flag bits
// (shengcompile.cpp:358-385, :447-505; struct sheng32 / sheng64).
int emit_sheng(const Dfa &d, const u8 *reach_map, const std::vector<u32> &program_of, u32 min_width, Buf &eng,
               u32 &state_bytes, std::string &err, bool wide = false) {
    const u32 n = (u32)d.next.size();
    const sheng::Shape sh = sheng::shape(!wide ? 17u : n <= 32 ? 21u : 22u);
    if (n > sh.max_states) { err = "DFA too large for a Sheng engine (" + std::to_string(n) + " states)"; return HSB_LIT_COMPILER_ERROR; }
    state_bytes = 1;
    bool can_die = false, any_eod = false;
    for (u32 i = 1; i < n; i++) for (u32 k = 0; k < d.n_sym; k++) if (d.next[i][k] == 0) can_die = true;
    auto state_byte = [&](u32 t) -> u8 {
        u8 v = (u8)t;
        if (!d.reports[t].empty()) v |= sh.accept;
        if (t == 0) v |= sh.dead;
        return v;
    };
    // report lists (normal and EOD share one pool, as gatherReports does)
    std::map<std::vector<u32>, u32> list_id;
    std::vector<std::vector<u32>> lists;
    std::set<u32> all_reports;
    auto intern = [&](const std::vector<u32> &pis) -> u32 {
        if (pis.empty()) return ~0u;
        std::set<u32> ps;
        for (u32 pi : pis) ps.insert(program_of[pi]);
        std::vector<u32> l(ps.begin(), ps.end());
        all_reports.insert(l.begin(), l.end());
        auto it = list_id.find(l);
        if (it == list_id.end()) { it = list_id.emplace(l, (u32)lists.size()).first; lists.push_back(l); }
        return it->second;
    };
    std::vector<u32> list_of(n), list_eod_of(n);
    for (u32 i = 0; i < n; i++) {
        list_of[i] = intern(d.reports[i]);
        list_eod_of[i] = intern(d.reports_eod[i]);
        if (list_eod_of[i] != ~0u) any_eod = true;
    }
    const size_t nfa_size = round_up(nfa::SIZE + sh.SIZE, 16);
    const size_t reports_offset = nfa_size + sheng::aux::SIZE * (size_t)n;
    size_t rl_size = 0;
    for (auto &l : lists) rl_size += 4 + 4 * l.size();
    const size_t accel_offset = round_up(reports_offset + rl_size, 16);
    const size_t total = round_up(accel_offset, 64);
    eng.grow(total);
    u8 *nh = eng.at(0), *m = eng.at(nfa::SIZE);
    wr32(nh, nfa::flags, any_eod ? nfa::ACCEPTS_EOD : 0);
    wr32(nh, nfa::length, (u32)total);
    wr8(nh, nfa::type, sh.nfa_type);
    wr32(nh, nfa::queueIndex, 0);
    wr32(nh, nfa::nPositions, n);
    wr32(nh, nfa::scratchStateSize, 1);
    wr32(nh, nfa::streamStateSize, 1);
    wr32(nh, nfa::maxWidth, 0);
    wr32(nh, nfa::minWidth, min_width);
    for (u32 c = 0; c < 256; c++)
        for (u32 i = 0; i < n; i++) {
            m[(size_t)sh.stride * c + i] = state_byte(d.next[i][reach_map[c]]);
            // Sheng32: the 64-byte permute looks at 6 bits of the state byte, one of which is its accept flag (0x20): every
            // entry is stored again 32 places up (createShuffleMasks<sheng32>, shengcompile.cpp:578-596)
            if (sh.nfa_type == 21) m[(size_t)sh.stride * c + 32 + i] = m[(size_t)sh.stride * c + i];
        }
    wr32(m, sh.length, (u32)(total - nfa::SIZE));
    wr32(m, sh.aux_offset, (u32)nfa_size);
    wr32(m, sh.report_offset, (u32)reports_offset);
    wr32(m, sh.accel_offset, (u32)accel_offset);
    wr8(m, sh.n_states, (u8)n);
    wr8(m, sh.anchored, state_byte(d.start_anchored));
    wr8(m, sh.floating, state_byte(d.start_floating));
    const bool single = all_reports.size() == 1;
    wr8(m, sh.flags, (u8)((single ? sheng::FLAG_SINGLE_REPORT : 0) | (can_die ? sheng::FLAG_CAN_DIE : 0)));
    wr32(m, sh.report, single ? *all_reports.begin() : 0);
    std::vector<u32> list_off(lists.size());
    size_t o = reports_offset;
    for (size_t i = 0; i < lists.size(); i++) {
        list_off[i] = (u32)o;
        wr32(eng.at(o), 0, (u32)lists[i].size());
        for (size_t k = 0; k < lists[i].size(); k++) wr32(eng.at(o), 4 + 4 * k, lists[i][k]);
        o += 4 + 4 * lists[i].size();
    }
    for (u32 i = 0; i < n; i++) {
        u8 *a = eng.at(nfa_size + sheng::aux::SIZE * (size_t)i);
        wr32(a, sheng::aux::accept, list_of[i] == ~0u ? 0 : list_off[list_of[i]]);
        wr32(a, sheng::aux::accept_eod, list_eod_of[i] == ~0u ? 0 : list_off[list_eod_of[i]]);
        wr32(a, sheng::aux::accel, 0);
        wr32(a, sheng::aux::top, state_byte(d.start_floating));
    }
    return HSB_LIT_SUCCESS;
}

}  // namespace

namespace {
using hsb_nfadb::Pattern;
struct Parsed { Glushkov g; Glushkov::Info info; bool anchored, eod; };

// parse every pattern of the set (the writer's dialect: classes, groups, alternation, * + ? {m,n}, ^, trailing \z)
int parse_all(const std::vector<Pattern> &pats, std::vector<Parsed> &parsed, std::string &err) {
    for (size_t pi = 0; pi < pats.size(); pi++) {
        const Pattern &p = pats[pi];
        if (p.flags & ~(HSB_FLAG_CASELESS | HSB_FLAG_DOTALL)) {
            err = "expression " + std::to_string(pi) + ": unsupported flag";
            return HSB_LIT_COMPILER_ERROR;
        }
        std::string body = p.expr;
        bool anchored = false;
        if (!body.empty() && body[0] == '^') { anchored = true; body.erase(0, 1); }
        // a trailing \\z: the match must end where the data ends (reported by the engine's EOD accepts)
        bool eod = false;
        if (body.size() >= 2 && body.compare(body.size() - 2, 2, "\\z") == 0 &&
            (body.size() < 3 || body[body.size() - 3] != '\\')) { eod = true; body.erase(body.size() - 2); }
        Parser ps(body, (p.flags & HSB_FLAG_CASELESS) != 0, (p.flags & HSB_FLAG_DOTALL) != 0);
        const int root = ps.parse_alt();
        if (root < 0 || !ps.eof()) {
            if (ps.err.empty()) ps.fail("unbalanced )");
            err = "expression " + std::to_string(pi) + ": " + ps.err;
            return HSB_LIT_COMPILER_ERROR;
        }
        if (ps.n_char_nodes > MAX_STATES) { err = "pattern set too large for one LimEx NFA"; return HSB_LIT_COMPILER_ERROR; }
        Parsed pr;
        pr.anchored = anchored;
        pr.eod = eod;
        pr.info = pr.g.build(ps.nodes, root);
        if (pr.info.nullable) {
            err = "expression " + std::to_string(pi) + ": matches the empty string (not supported)";
            return HSB_LIT_COMPILER_ERROR;
        }
        parsed.push_back(std::move(pr));
    }
    return HSB_LIT_SUCCESS;
}

// one Glushkov automaton for the patterns `members` (indices into parsed); report_of holds GLOBAL pattern index + 1.
// Returns false when the automaton needs more than MAX_STATES states.
bool make_automaton(const std::vector<Parsed> &parsed, const std::vector<u32> &members, Automaton &am, size_t &total) {
    bool any_unanch = false, any_anch = false;
    for (u32 pi : members) (parsed[pi].anchored ? any_anch : any_unanch) = true;
    total = (any_unanch ? 1 : 0) + (any_anch ? 1 : 0);
    for (u32 pi : members) total += parsed[pi].g.cls.size();
    if (total > MAX_STATES) return false;
    am = Automaton();
    am.n_states = (u32)total;
    am.reach_of.assign(total, CharSet());
    am.succ.assign(total, StateSet());
    am.report_of.assign(total, 0);
    am.report_eod_of.assign(total, 0);
    u32 next = 0;
    const u32 st_ds = any_unanch ? next++ : ~0u, st_anch = any_anch ? next++ : ~0u;
    if (any_unanch) {
        am.reach_of[st_ds].invert();  // entered on every byte
        am.succ[st_ds].set(st_ds);
        am.init.set(st_ds);
        am.init_ds.set(st_ds);
    }
    if (any_anch) am.init.set(st_anch);
    for (u32 pi : members) {
        const Parsed &pr = parsed[pi];
        const u32 base = next;
        const u32 n = (u32)pr.g.cls.size();
        for (u32 p = 0; p < n; p++) {
            am.reach_of[base + p] = pr.g.cls[p];
            for (u32 q = 0; q < n; q++) if (pr.g.follow[p].test(q)) am.succ[base + p].set(base + q);
            if (pr.info.last.test(p)) (pr.eod ? am.report_eod_of : am.report_of)[base + p] = pi + 1;
            if (pr.eod) am.any_eod = true;
            if (pr.info.first.test(p)) am.succ[pr.anchored ? st_anch : st_ds].set(base + p);
        }
        am.min_width = std::min(am.min_width, pr.info.min_len);
        next += n;
    }
    return true;
}

// reach table: distinct per-byte masks
bool reach_table(const Automaton &am, std::vector<StateSet> &reach_rows, u8 *reach_map) {
    reach_rows.clear();
    std::map<std::string, u32> seen;
    for (u32 c = 0; c < 256; c++) {
        StateSet row;
        for (u32 s = 0; s < am.n_states; s++) if (am.reach_of[s].test(c)) row.set(s);
        const std::string key = row.to_string();
        auto it = seen.find(key);
        if (it == seen.end()) { it = seen.emplace(key, (u32)reach_rows.size()).first; reach_rows.push_back(row); }
        reach_map[c] = (u8)it->second;
    }
    return reach_rows.size() <= 256;
}

// LimEx model (shifts + exceptions) and layout of one automaton: struct NFA + LimExNFA<bits> + tables, relative to the
// engine's own start.  program_of[global pattern index] = Rose report program offset.
int emit_limex(const Automaton &am, const std::vector<StateSet> &reach_rows, const u8 *reach_map,
               const std::vector<u32> &program_of, Buf &eng, u32 &stream_state, u32 &scratch_state, std::string &err) {
    (void)err;
    const size_t n_pats = program_of.size();
    u32 type = nfa::LIMEX_512;
    for (u32 t = 0; t < 6; t++) if (limex::layout(t).bits >= am.n_states) { type = t; break; }
    const limex::Layout L = limex::layout(type);
    // most frequent forward distances become shifts
    std::map<u32, u32> dist_count;
    for (u32 s = 0; s < am.n_states; s++)
        for (u32 d = 0; d <= limex::MAX_SHIFT_AMOUNT && s + d < am.n_states; d++)
            if (am.succ[s].test(s + d) && (s >> 6) == ((s + d) >> 6)) dist_count[d]++;
    std::vector<std::pair<u32, u32>> by_count;  // (count, distance)
    for (auto &dc : dist_count) by_count.push_back({dc.second, dc.first});
    std::sort(by_count.begin(), by_count.end(), [](auto &a, auto &b) { return a.first != b.first ? a.first > b.first : a.second < b.second; });
    std::vector<u32> shifts;
    for (auto &c : by_count) if (shifts.size() < limex::MAX_SHIFT_COUNT) shifts.push_back(c.second);
    if (shifts.empty()) shifts.push_back(1);  // "number of shifts should always be >= 1"
    std::sort(shifts.begin(), shifts.end());
    std::vector<StateSet> shift_mask(shifts.size());
    std::vector<StateSet> exc_succ(am.n_states);
    StateSet exception_mask, accept_mask, accept_eod_mask;
    for (u32 s = 0; s < am.n_states; s++) {
        for (u32 t = 0; t < am.n_states; t++) {
            if (!am.succ[s].test(t)) continue;
            bool shifted = false;
            // a shift never crosses a 64-bit lane of the state vector (limex_compile.cpp:245-255)
            if (t >= s && (s >> 6) == (t >> 6))
                for (size_t k = 0; k < shifts.size(); k++)
                    if (shifts[k] == t - s) { shift_mask[k].set(s); shifted = true; }
            if (!shifted) { exc_succ[s].set(t); exception_mask.set(s); }
        }
        if (am.report_of[s]) { exception_mask.set(s); accept_mask.set(s); }
        if (am.report_eod_of[s]) accept_eod_mask.set(s);  // EOD accepts fire from the accept table only
    }
    eng.grow(nfa::SIZE + L.SIZE);
    const size_t lx = nfa::SIZE;  // LimExNFA base inside eng
    // reach table right behind the struct (get_reach_table)
    const size_t reach_off = eng.grow(reach_rows.size() * L.bytes);
    for (size_t i = 0; i < reach_rows.size(); i++) put_state(eng.at(reach_off + i * L.bytes), reach_rows[i], L.bytes);
    eng.align(L.align);
    const size_t top_off = eng.grow(L.bytes);
    put_state(eng.at(top_off), am.init, L.bytes);
    const size_t accel_tab_off = eng.grow(1);  // one entry: state pattern 0 -> aux 0
    eng.align(16);
    const size_t accel_aux_off = eng.grow(hwlm::ACCEL_AUX_SIZE);  // aux[0] = ACCEL_NONE
    eng.align(4);
    std::vector<u32> accept_states, accept_eod_states, exc_states;
    for (u32 s = 0; s < am.n_states; s++) {
        if (accept_mask.test(s)) accept_states.push_back(s);
        if (accept_eod_mask.test(s)) accept_eod_states.push_back(s);
        if (exception_mask.test(s)) exc_states.push_back(s);
    }
    const size_t accept_off = eng.grow(accept_states.size() * nfaaccept::SIZE);
    for (size_t i = 0; i < accept_states.size(); i++) {
        u8 *a = eng.at(accept_off + i * nfaaccept::SIZE);
        wr8(a, nfaaccept::single_report, 1);
        wr32(a, nfaaccept::reports, program_of[am.report_of[accept_states[i]] - 1]);
        wr32(a, nfaaccept::squash, 0xffffffffu);
    }
    const size_t accept_eod_off = eng.grow(accept_eod_states.size() * nfaaccept::SIZE);
    for (size_t i = 0; i < accept_eod_states.size(); i++) {
        u8 *a = eng.at(accept_eod_off + i * nfaaccept::SIZE);
        wr8(a, nfaaccept::single_report, 1);
        wr32(a, nfaaccept::reports, program_of[am.report_eod_of[accept_eod_states[i]] - 1]);
        wr32(a, nfaaccept::squash, 0xffffffffu);
    }
    // report lists for the exceptions: {program offset, MO_INVALID_IDX} per pattern
    eng.align(4);
    std::vector<u32> report_list_of(n_pats);
    for (size_t pi = 0; pi < n_pats; pi++) {
        const size_t o = eng.grow(8);
        wr32(eng.at(o), 0, program_of[pi]);
        wr32(eng.at(o), 4, 0xffffffffu);
        report_list_of[pi] = (u32)(o - lx);
    }
    eng.align(L.align < 16 ? 16 : L.align);
    const size_t exc_off = eng.grow(exc_states.size() * L.E_SIZE);
    for (size_t i = 0; i < exc_states.size(); i++) {
        u8 *e = eng.at(exc_off + i * L.E_SIZE);
        const u32 s = exc_states[i];
        StateSet ones;
        ones.set();
        put_state(e + L.e_squash, ones, L.bytes);
        put_state(e + L.e_successors, exc_succ[s], L.bytes);
        wr32(e, L.e_reports, am.report_of[s] ? report_list_of[am.report_of[s] - 1] : 0xffffffffu);
        wr32(e, L.e_repeatOffset, 0xffffffffu);
        wr8(e, L.e_hasSquash, limex::SQUASH_NONE);
        wr8(e, L.e_trigger, 0);
    }
    eng.align(CACHELINE);
    // struct NFA
    u8 *nh = eng.at(0);
    wr32(nh, nfa::flags, am.any_eod ? nfa::ACCEPTS_EOD : 0);
    wr32(nh, nfa::length, (u32)eng.size());
    wr8(nh, nfa::type, (u8)type);
    wr32(nh, nfa::queueIndex, 0);
    wr32(nh, nfa::nPositions, am.n_states);
    wr32(nh, nfa::scratchStateSize, L.bytes);
    stream_state = (am.n_states + 7) / 8;
    scratch_state = L.bytes;
    wr32(nh, nfa::streamStateSize, stream_state);
    wr32(nh, nfa::maxWidth, 0);
    wr32(nh, nfa::minWidth, am.min_width);
    // struct LimExNFA
    u8 *lp = eng.at(lx);
    std::memcpy(lp + limex::reachMap, reach_map, 256);
    wr32(lp, limex::reachSize, (u32)reach_rows.size());
    wr32(lp, limex::accelCount, 0);
    wr32(lp, limex::accelTableOffset, (u32)(accel_tab_off - lx));
    wr32(lp, limex::accelAuxCount, 1);
    wr32(lp, limex::accelAuxOffset, (u32)(accel_aux_off - lx));
    wr32(lp, limex::acceptCount, (u32)accept_states.size());
    wr32(lp, limex::acceptOffset, (u32)(accept_off - lx));
    wr32(lp, limex::acceptEodCount, (u32)accept_eod_states.size());
    wr32(lp, limex::acceptEodOffset, (u32)(accept_eod_off - lx));
    wr32(lp, limex::exceptionCount, (u32)exc_states.size());
    wr32(lp, limex::exceptionOffset, (u32)(exc_off - lx));
    wr32(lp, limex::repeatCount, 0);
    wr32(lp, limex::repeatOffset, 0);
    wr32(lp, limex::squashOffset, 0);
    wr32(lp, limex::squashCount, 0);
    wr32(lp, limex::topCount, 1);
    wr32(lp, limex::topOffset, (u32)(top_off - lx));
    wr32(lp, limex::stateSize, stream_state);
    wr32(lp, limex::flags, 0);
    StateSet all;
    for (u32 s = 0; s < am.n_states; s++) all.set(s);
    put_state(lp + L.init, am.init, L.bytes);
    put_state(lp + L.initDS, am.init_ds, L.bytes);
    put_state(lp + L.accept, accept_mask, L.bytes);
    put_state(lp + L.acceptAtEOD, accept_eod_mask, L.bytes);
    put_state(lp + L.compressMask, all, L.bytes);
    put_state(lp + L.exceptionMask, exception_mask, L.bytes);
    for (size_t k = 0; k < shifts.size(); k++) {
        put_state(lp + L.shift + k * L.bytes, shift_mask[k], L.bytes);
        wr8(lp, L.shiftAmount + k, (u8)shifts[k]);
    }
    wr32(lp, L.shiftCount, (u32)shifts.size());
    return HSB_LIT_SUCCESS;
}

// one engine for one automaton: LimEx, or the determinised McClellan / Sheng form
int emit_engine(const Automaton &am, unsigned engine, const std::vector<u32> &program_of, u32 dfa_state_limit, Buf &eng,
                u32 &stream_state, u32 &scratch_state, std::string &err, bool sheng_if_small = false) {
    std::vector<StateSet> reach_rows;
    u8 reach_map[256];
    if (!reach_table(am, reach_rows, reach_map)) { err = "internal: reach table overflow"; return HSB_LIT_COMPILER_ERROR; }
    if (engine == HSB_REGEX_ENGINE_LIMEX) return emit_limex(am, reach_rows, reach_map, program_of, eng, stream_state, scratch_state, err);
    Dfa dfa;
    if (!determinise(am, reach_rows, dfa, dfa_state_limit)) { err = "DFA too large for a McClellan engine"; return HSB_LIT_COMPILER_ERROR; }
    minimise(dfa);
    // outfix engines of multi-engine databases: a DFA of at most 16 states is laid out as a Sheng engine, which is what the
    // reference compiler tries first for every DFA (shengCompile before mcclellanCompile, src/rose/rose_build_bytecode.cpp
    // buildOutfix / getDfa)
    if (sheng_if_small && (engine == HSB_REGEX_ENGINE_MCCLELLAN || engine == HSB_REGEX_ENGINE_MCSHENG || engine == HSB_REGEX_ENGINE_MCCLELLAN_WIDE || engine == HSB_REGEX_ENGINE_MCSHENG64) && dfa.next.size() <= 16)
        engine = HSB_REGEX_ENGINE_SHENG;
    if (engine == HSB_REGEX_ENGINE_MCSHENG || engine == HSB_REGEX_ENGINE_MCSHENG64) {
        const int mrv = emit_mcsheng(dfa, reach_map, program_of, am.min_width, eng, stream_state, err, engine == HSB_REGEX_ENGINE_MCSHENG64);
        scratch_state = stream_state;
        if (mrv == HSB_LIT_SUCCESS || !sheng_if_small) return mrv;
        engine = HSB_REGEX_ENGINE_MCCLELLAN;  // no region (the start state accepts): plain McClellan, as getDfa falls through
        err.clear();
    }
    const int rv = engine == HSB_REGEX_ENGINE_SHENG || engine == HSB_REGEX_ENGINE_SHENG_WIDE
                       ? emit_sheng(dfa, reach_map, program_of, am.min_width, eng, stream_state, err, engine == HSB_REGEX_ENGINE_SHENG_WIDE)
                       : emit_mcclellan(dfa, reach_map, program_of, engine == HSB_REGEX_ENGINE_MCCLELLAN_SHERMAN || engine == HSB_REGEX_ENGINE_MCCLELLAN_WIDE,
                                        am.min_width, eng, stream_state, err, engine == HSB_REGEX_ENGINE_MCCLELLAN_WIDE);
    scratch_state = stream_state;
    return rv;
}

}  // namespace

namespace hsb_nfadb {

int build(const std::vector<Pattern> &pats, unsigned engine, std::vector<u8> &image, std::string &err, unsigned mode) {
    if (engine > HSB_REGEX_ENGINE_MCSHENG64) { err = "unknown engine"; return HSB_LIT_COMPILER_ERROR; }
    if (pats.empty()) { err = "no expressions"; return HSB_LIT_COMPILER_ERROR; }
    // ---- 1. parse every pattern, build one automaton -------------------------------------------------
    std::vector<Parsed> parsed;
    if (parse_all(pats, parsed, err) != HSB_LIT_SUCCESS) return HSB_LIT_COMPILER_ERROR;
    Automaton am;
    std::vector<u32> everyone(pats.size());
    for (u32 i = 0; i < pats.size(); i++) everyone[i] = i;
    size_t total = 0;
    if (!make_automaton(parsed, everyone, am, total)) { err = "pattern set too large for one LimEx NFA (" + std::to_string(total) + " states)"; return HSB_LIT_COMPILER_ERROR; }

    // ---- 3. Rose blob: header, dedupe keys, report programs, NfaInfo, NFA ------------------------------
    Buf rose;
    rose.grow(rose::SIZE);
    rose.align(CACHELINE);
    // one dedupe key per distinct report id: two positions of a pattern may accept at the same offset
    std::map<u32, u32> dkeys;
    for (auto &p : pats) if (!dkeys.count(p.id)) { u32 k = (u32)dkeys.size(); dkeys[p.id] = k; }
    const u32 inv_dkey_off = (u32)rose.grow(4 * dkeys.size());
    for (auto &d : dkeys) wr32(rose.at(inv_dkey_off), 4 * d.second, d.first);
    rose.align(op::ALIGN);
    std::vector<u32> program_of(pats.size());
    for (size_t pi = 0; pi < pats.size(); pi++) {
        rose.align(op::ALIGN);
        const u32 o = (u32)rose.grow(op::dedupe_and_report::SIZE + op::SIZE_1);
        program_of[pi] = o;
        wr8(rose.at(o), 0, op::DEDUPE_AND_REPORT);
        wr8(rose.at(o), op::dedupe_and_report::quash_som, 0);
        wr32(rose.at(o), op::dedupe_and_report::dkey, dkeys[pats[pi].id]);
        wr32(rose.at(o), op::dedupe_and_report::onmatch, pats[pi].id);
        wr32(rose.at(o), op::dedupe_and_report::offset_adjust, 0);
        wr32(rose.at(o), op::dedupe_and_report::fail_jump, (u32)op::dedupe_and_report::SIZE);
        wr8(rose.at(o + op::dedupe_and_report::SIZE), 0, op::END);
    }
    rose.align(4);
    const u32 nfa_info_off = (u32)rose.grow(nfainfo::SIZE);
    rose.align(CACHELINE);
    const u32 nfa_off = (u32)rose.size();

    // the engine, relative to its own start
    Buf eng;
    u32 stream_state = 0, scratch_state = 0;
    if (emit_engine(am, engine, program_of, 4 * (mcclellan::STATE_MASK + 1), eng, stream_state, scratch_state, err) != HSB_LIT_SUCCESS)
        return HSB_LIT_COMPILER_ERROR;

    const size_t nfa_at = rose.grow(eng.size());
    std::memcpy(rose.at(nfa_at), eng.b.data(), eng.size());
    rose.align(CACHELINE);

    // NfaInfo of queue 0
    u8 *r = rose.at(0);
    // stream-state layout (fillStateOffsets): status byte, leaf multibit of one engine, no groups / history /
    // exhaustion, then the NFA's stream state
    u32 cur = 1;
    u8 *so = r + rose::stateOffsets;
    wr32(so, rose::so::activeLeafArray, cur);
    wr32(so, rose::so::activeLeafArray_size, mmbit_size(1));
    cur += mmbit_size(1);
    wr32(so, rose::so::activeLeftArray, cur);
    wr32(so, rose::so::longLitState, cur);
    wr32(so, rose::so::leftfixLagTable, cur);
    wr32(so, rose::so::anchorState, cur);
    wr32(so, rose::so::groups, cur);
    wr32(so, rose::so::groups_size, 0);
    wr32(so, rose::so::history, cur);
    wr32(so, rose::so::exhausted, cur);
    wr32(so, rose::so::exhausted_size, mmbit_size(0));
    wr32(so, rose::so::logicalVec, cur);
    wr32(so, rose::so::combVec, cur);
    wr32(so, rose::so::nfaStateBegin, cur);
    const u32 nfa_stream_off = cur;
    cur += stream_state;
    wr32(so, rose::so::end, cur);
    u8 *ni = rose.at(nfa_info_off);
    wr32(ni, nfainfo::nfaOffset, nfa_off);
    wr32(ni, nfainfo::stateOffset, nfa_stream_off);
    wr32(ni, nfainfo::fullStateOffset, 0);

    wr8(r, rose::pureLiteral, 0);
    wr8(r, rose::runtimeImpl, rose::RUNTIME_SINGLE_OUTFIX);
    wr8(r, rose::canExhaust, 0);
    wr32(r, rose::mode, mode);   // stream mode: the engine's stream state sits at NfaInfo.stateOffset of the stream state
    // hs_close_stream runs the engine's end-of-data check only when the header asks for it (report_eod_matches,
    // src/runtime.c:676-692 -> soleOutfixEodExec); the block path looks at the engine's own flag instead (:273-278)
    if (mode == MODE_STREAM && am.any_eod) wr8(r, rose::requiresEodCheck, 1);
    wr32(r, rose::ekeyCount, 0);
    wr32(r, rose::dkeyCount, (u32)dkeys.size());
    wr32(r, rose::dkeyLogSize, fatbit_size((u32)dkeys.size()));
    wr32(r, rose::invDkeyOffset, inv_dkey_off);
    wr32(r, rose::somLocationFatbitSize, fatbit_size(0));
    wr32(r, rose::stateSize, cur);
    wr32(r, rose::scratchStateSize, (u32)round_up(scratch_state, 64));
    wr32(r, rose::amatcherMaxBiAnchoredWidth, rose::BOUND_INF);
    wr32(r, rose::fmatcherMaxBiAnchoredWidth, rose::BOUND_INF);
    wr32(r, rose::activeArrayCount, 1);
    wr32(r, rose::queueCount, 1);
    wr32(r, rose::activeQueueArraySize, fatbit_size(1));
    wr32(r, rose::handledKeyFatbitSize, fatbit_size(0));
    wr32(r, rose::minWidth, am.min_width);
    wr32(r, rose::minWidthExcludingBoundaries, am.min_width);
    wr32(r, rose::maxBiAnchoredWidth, rose::BOUND_INF);
    wr32(r, rose::nfaInfoOffset, nfa_info_off);
    wr64(r, rose::initialGroups, 0);
    wr64(r, rose::floating_group_mask, 0);
    wr32(r, rose::size, (u32)rose.size());
    wr32(r, rose::delay_fatbit_size, fatbit_size(0));
    wr32(r, rose::anchored_fatbit_size, fatbit_size(0));
    wr32(r, rose::outfixBeginQueue, 0);
    wr32(r, rose::outfixEndQueue, 1);
    wr32(r, rose::leftfixBeginQueue, 1);
    wr32(r, rose::initMpvNfa, 0xffffffffu);

    const size_t bc_off = db::IMAGE_BYTECODE;
    image.assign(db::SIZE + rose.size(), 0);
    std::memcpy(image.data() + bc_off, rose.b.data(), rose.size());
    wr32(image.data(), db::magic, db::MAGIC);
    wr32(image.data(), db::version, db::VERSION_5_4_11);
    wr32(image.data(), db::length, (u32)rose.size());
    wr64(image.data(), db::platform, db::PLATFORM_GENERIC);
    wr32(image.data(), db::crc32, hsb_crc32c(rose.b.data(), rose.size()));
    wr32(image.data(), db::bytecode, (u32)bc_off);
    return HSB_LIT_SUCCESS;
}


// ---- full-Rose database of outfix engines only ------------------------------------------------------------------
// The shape the reference compiler gives a pattern set without usable literals: runtimeImpl = ROSE_RUNTIME_FULL_ROSE, no
// literal matchers, no roles; every group of patterns is one OUTFIX engine on its own queue [outfixBeginQueue,
// outfixEndQueue).  At scan time roseBlockExec (src/rose/block.c:345-422) starts every outfix (blockInitSufPQ,
// src/rose/catchup.c:612-672: MQE_START, MQE_TOP, MQE_END), and roseCatchUpTo (catchup.h:128-167, roseCatchUpAll
// catchup.c:560+) interleaves their matches by offset through the report programs.  Followed for the layout:
// rose_build_bytecode.cpp:2078-2116 (state space per engine), :3608-3890 (header), rose_build_scatter.cpp:66-97
// (state_init plan: the active leaf array is cleared before every block).
// Grouping: patterns join the current group while its determinised automaton stays within `dfa_states` states; a pattern
// whose own DFA is larger goes to a LimEx group (<= 512 NFA states) -- the reference's own fallback when McClellan
// construction fails (src/nfagraph/ng_limex.cpp / ng_mcclellan.cpp state limits).
int build_outfix_engines(const std::vector<Pattern> &pats, unsigned engine, u32 dfa_states,
                         const std::vector<u32> &program_of, std::vector<EngineBlob> &out, std::string &err) {
    if (engine != HSB_REGEX_ENGINE_LIMEX && engine != HSB_REGEX_ENGINE_MCCLELLAN && engine != HSB_REGEX_ENGINE_MCCLELLAN_SHERMAN &&
        engine != HSB_REGEX_ENGINE_MCSHENG && engine != HSB_REGEX_ENGINE_MCCLELLAN_WIDE && engine != HSB_REGEX_ENGINE_MCSHENG64) {
        err = "outfix engines are LimEx, McClellan or McSheng";
        return HSB_LIT_COMPILER_ERROR;
    }
    if (pats.empty()) { err = "no expressions"; return HSB_LIT_COMPILER_ERROR; }
    if (!dfa_states) dfa_states = 1024;
    if (dfa_states > mcclellan::STATE_MASK) dfa_states = mcclellan::STATE_MASK;
    std::vector<Parsed> parsed;
    if (parse_all(pats, parsed, err) != HSB_LIT_SUCCESS) return HSB_LIT_COMPILER_ERROR;
    struct Group { std::vector<u32> members; unsigned engine; };
    std::vector<Group> groups;
    auto dfa_fits = [&](const std::vector<u32> &members) {
        Automaton am;
        size_t total = 0;
        if (!make_automaton(parsed, members, am, total)) return false;
        std::vector<StateSet> rows;
        u8 map[256];
        if (!reach_table(am, rows, map)) return false;
        Dfa d;
        return determinise(am, rows, d, dfa_states);
    };
    std::vector<u32> for_limex;
    if (engine == HSB_REGEX_ENGINE_LIMEX) {
        for (u32 i = 0; i < pats.size(); i++) for_limex.push_back(i);
    } else {
        // patterns are visited in the order of their text (anchored ones first), so that patterns sharing a head --
        // which share DFA states -- are neighbours; identical expressions cost no states at all
        std::vector<u32> order(pats.size());
        for (u32 i = 0; i < pats.size(); i++) order[i] = i;
        std::stable_sort(order.begin(), order.end(), [&](u32 a, u32 b) {
            if (parsed[a].anchored != parsed[b].anchored) return parsed[a].anchored;
            return pats[a].expr < pats[b].expr;
        });
        std::vector<u32> cur;
        for (u32 i : order) {
            cur.push_back(i);
            if (dfa_fits(cur)) continue;
            cur.pop_back();
            if (!cur.empty()) { groups.push_back({cur, engine}); cur.clear(); }
            cur.push_back(i);
            if (!dfa_fits(cur)) { cur.clear(); for_limex.push_back(i); }
        }
        if (!cur.empty()) groups.push_back({cur, engine});
    }
    {
        std::vector<u32> cur;
        size_t states = 2;
        for (u32 i : for_limex) {
            const size_t n = parsed[i].g.cls.size();
            if (n + 2 > MAX_STATES) { err = "expression " + std::to_string(i) + " needs more than 512 NFA states"; return HSB_LIT_COMPILER_ERROR; }
            if (states + n > MAX_STATES) { groups.push_back({cur, HSB_REGEX_ENGINE_LIMEX}); cur.clear(); states = 2; }
            cur.push_back(i);
            states += n;
        }
        if (!cur.empty()) groups.push_back({cur, HSB_REGEX_ENGINE_LIMEX});
    }
    if (groups.size() > 4096) { err = "more than 4096 engines"; return HSB_LIT_COMPILER_ERROR; }
    for (u32 qi = 0; qi < groups.size(); qi++) {
        Automaton am;
        size_t total = 0;
        if (!make_automaton(parsed, groups[qi].members, am, total)) { err = "internal: group too large"; return HSB_LIT_COMPILER_ERROR; }
        Buf eng;
        EngineBlob blob;
        if (emit_engine(am, groups[qi].engine, program_of, 4 * (mcclellan::STATE_MASK + 1), eng, blob.stream_state, blob.scratch_state, err, true) != HSB_LIT_SUCCESS)
            return HSB_LIT_COMPILER_ERROR;
        wr32(eng.at(0), nfa::queueIndex, qi);
        blob.bytes = eng.b;
        blob.min_width = am.min_width;
        out.push_back(std::move(blob));
    }
    return HSB_LIT_SUCCESS;
}

int build_rose(const std::vector<Pattern> &pats, unsigned engine, u32 dfa_states, std::vector<u8> &image, std::string &err) {
    if (pats.empty()) { err = "no expressions"; return HSB_LIT_COMPILER_ERROR; }
    Buf rose;
    rose.grow(rose::SIZE);
    rose.align(CACHELINE);
    std::map<u32, u32> dkeys;
    for (auto &p : pats) if (!dkeys.count(p.id)) { u32 k = (u32)dkeys.size(); dkeys[p.id] = k; }
    const u32 inv_dkey_off = (u32)rose.grow(4 * dkeys.size());
    for (auto &d : dkeys) wr32(rose.at(inv_dkey_off), 4 * d.second, d.first);
    rose.align(op::ALIGN);
    std::vector<u32> program_of(pats.size());
    for (size_t pi = 0; pi < pats.size(); pi++) {
        rose.align(op::ALIGN);
        const u32 o = (u32)rose.grow(op::dedupe_and_report::SIZE + op::SIZE_1);
        program_of[pi] = o;
        wr8(rose.at(o), 0, op::DEDUPE_AND_REPORT);
        wr32(rose.at(o), op::dedupe_and_report::dkey, dkeys[pats[pi].id]);
        wr32(rose.at(o), op::dedupe_and_report::onmatch, pats[pi].id);
        wr32(rose.at(o), op::dedupe_and_report::fail_jump, (u32)op::dedupe_and_report::SIZE);
        wr8(rose.at(o + op::dedupe_and_report::SIZE), 0, op::END);
    }
    std::vector<EngineBlob> blobs;
    if (build_outfix_engines(pats, engine, dfa_states, program_of, blobs, err) != HSB_LIT_SUCCESS) return HSB_LIT_COMPILER_ERROR;
    const u32 N = (u32)blobs.size();
    rose.align(4);
    const u32 nfa_info_off = (u32)rose.grow((size_t)nfainfo::SIZE * N);
    // state layout: status byte, active leaf multibit of N queues, then every engine's stream state
    const u32 leaf_off = 1, leaf_bytes = mmbit_size(N);
    u32 cur_state = leaf_off + leaf_bytes;
    const u32 nfa_state_begin = cur_state;
    u32 scratch_total = 0, min_width = ~0u;
    for (u32 qi = 0; qi < N; qi++) {
        const EngineBlob &blob = blobs[qi];
        rose.align(CACHELINE);
        const u32 nfa_off = (u32)rose.grow(blob.bytes.size());
        std::memcpy(rose.at(nfa_off), blob.bytes.data(), blob.bytes.size());
        u8 *ni = rose.at(nfa_info_off + (size_t)nfainfo::SIZE * qi);
        wr32(ni, nfainfo::nfaOffset, nfa_off);
        wr32(ni, nfainfo::stateOffset, cur_state);
        scratch_total = (u32)round_up(scratch_total, 64);
        wr32(ni, nfainfo::fullStateOffset, scratch_total);
        cur_state += blob.stream_state;
        scratch_total += blob.scratch_state;
        min_width = std::min(min_width, blob.min_width);
    }
    const u32 eod_program = emit_engines_eod(rose, blobs, err);
    if (eod_program == ~0u) return HSB_LIT_COMPILER_ERROR;
    for (u32 qi = 0; qi < N; qi++) wr8(rose.at(nfa_info_off + (size_t)nfainfo::SIZE * qi), nfainfo::eod, accepts_eod(blobs[qi]) ? 1 : 0);
    rose.align(CACHELINE);
    // state_init: clear the active leaf array (8-byte stores, then single bytes)
    u32 plan_u64_off = 0, plan_u64_n = 0, plan_u8_off = 0, plan_u8_n = 0;
    {
        u32 b = 0;
        if (N > 256) {
            // hierarchical multibit: clearing the root block empties it (mmbit_clear, src/util/multibit.h:332-343)
            plan_u64_off = (u32)rose.size();
            const size_t o = rose.grow(16);
            wr32(rose.at(o), 0, leaf_off);
            plan_u64_n = 1;
            b = leaf_bytes;
        } else if (leaf_bytes >= 8) {
            plan_u64_off = (u32)rose.size();
            for (; b + 8 <= leaf_bytes; b += 8) {
                const size_t o = rose.grow(16);  // struct scatter_unit_u64a { u32 offset; u64a val; }
                wr32(rose.at(o), 0, leaf_off + b);
                plan_u64_n++;
            }
        }
        if (b < leaf_bytes) {
            plan_u8_off = (u32)rose.size();
            for (; b < leaf_bytes; b++) {
                const size_t o = rose.grow(8);   // struct scatter_unit_u8 { u32 offset; u8 val; }
                wr32(rose.at(o), 0, leaf_off + b);
                plan_u8_n++;
            }
        }
    }
    rose.align(CACHELINE);

    u8 *r = rose.at(0);
    u8 *so = r + rose::stateOffsets;
    wr32(so, rose::so::activeLeafArray, leaf_off);
    wr32(so, rose::so::activeLeafArray_size, leaf_bytes);
    for (size_t f : {rose::so::activeLeftArray, rose::so::longLitState, rose::so::leftfixLagTable, rose::so::anchorState,
                     rose::so::groups, rose::so::history, rose::so::exhausted, rose::so::logicalVec, rose::so::combVec,
                     rose::so::nfaStateBegin})
        wr32(so, f, nfa_state_begin);
    wr32(so, rose::so::exhausted_size, mmbit_size(0));
    wr32(so, rose::so::end, cur_state);
    // struct scatter_full_plan (src/util/scatter.h:43-52)
    u8 *sp = r + rose::state_init;
    wr32(sp, 0, plan_u64_off); wr32(sp, 4, plan_u64_n);
    wr32(sp, 24, plan_u8_n); wr32(sp, 28, plan_u8_off);

    wr8(r, rose::pureLiteral, 0);
    wr8(r, rose::runtimeImpl, rose::RUNTIME_FULL_ROSE);
    wr8(r, rose::hasOutfixesInSmallBlock, 1);
    wr8(r, rose::requiresEodCheck, eod_program ? 1 : 0);
    wr32(r, rose::eodProgramOffset, eod_program);
    wr32(r, rose::mode, MODE_BLOCK);
    wr32(r, rose::dkeyCount, (u32)dkeys.size());
    wr32(r, rose::dkeyLogSize, fatbit_size((u32)dkeys.size()));
    wr32(r, rose::invDkeyOffset, inv_dkey_off);
    wr32(r, rose::somLocationFatbitSize, fatbit_size(0));
    wr32(r, rose::stateSize, cur_state);
    wr32(r, rose::scratchStateSize, (u32)round_up(scratch_total, 64));
    wr32(r, rose::amatcherMaxBiAnchoredWidth, rose::BOUND_INF);
    wr32(r, rose::fmatcherMaxBiAnchoredWidth, rose::BOUND_INF);
    wr32(r, rose::activeArrayCount, N);
    wr32(r, rose::queueCount, N);
    wr32(r, rose::activeQueueArraySize, fatbit_size(N));
    wr32(r, rose::handledKeyFatbitSize, fatbit_size(0));
    wr32(r, rose::minWidth, min_width);
    wr32(r, rose::minWidthExcludingBoundaries, min_width);
    wr32(r, rose::maxBiAnchoredWidth, rose::BOUND_INF);
    wr32(r, rose::floatingDistance, rose::BOUND_INF);
    wr32(r, rose::nfaInfoOffset, nfa_info_off);
    wr32(r, rose::size, (u32)rose.size());
    wr32(r, rose::delay_fatbit_size, fatbit_size(0));
    wr32(r, rose::anchored_fatbit_size, fatbit_size(0));
    wr32(r, rose::outfixBeginQueue, 0);
    wr32(r, rose::outfixEndQueue, N);
    wr32(r, rose::leftfixBeginQueue, N);
    wr32(r, rose::initMpvNfa, 0xffffffffu);

    const size_t bc_off = db::IMAGE_BYTECODE;
    image.assign(db::SIZE + rose.size(), 0);
    std::memcpy(image.data() + bc_off, rose.b.data(), rose.size());
    wr32(image.data(), db::magic, db::MAGIC);
    wr32(image.data(), db::version, db::VERSION_5_4_11);
    wr32(image.data(), db::length, (u32)rose.size());
    wr64(image.data(), db::platform, db::PLATFORM_GENERIC);
    wr32(image.data(), db::crc32, hsb_crc32c(rose.b.data(), rose.size()));
    wr32(image.data(), db::bytecode, (u32)bc_off);
    return HSB_LIT_SUCCESS;
}

}  // namespace hsb_nfadb

extern "C" int hsb_compile_regex_multi_ex(const char *const *expressions, const unsigned *flags, const unsigned *ids,
                                          unsigned elements, unsigned mode, unsigned engine, hs_database_t **db,
                                          char **error) {
    if (error) *error = nullptr;
    if (!db) return HSB_LIT_COMPILER_ERROR;
    *db = nullptr;
    auto fail = [&](const std::string &m) {
        if (error) {
            char *p = (char *)std::malloc(m.size() + 1);
            if (p) std::memcpy(p, m.c_str(), m.size() + 1);
            *error = p;
        }
        return HSB_LIT_COMPILER_ERROR;
    };
    if (!expressions || !elements) return fail("invalid parameter: no expressions");
    if (mode != hsb::MODE_BLOCK && mode != hsb::MODE_STREAM) return fail("only HS_MODE_BLOCK and HS_MODE_STREAM databases are written");
    if (mode == hsb::MODE_STREAM && engine == HSB_REGEX_ENGINE_LIMEX) return fail("HS_MODE_STREAM is written for DFA engines only");
    std::vector<hsb_nfadb::Pattern> pats;
    for (unsigned i = 0; i < elements; i++) {
        if (!expressions[i]) return fail("null expression");
        pats.push_back({expressions[i], flags ? flags[i] : 0, ids ? ids[i] : 0});
    }
    std::vector<hsb::u8> image;
    std::string err;
    if (hsb_nfadb::build(pats, engine, image, err, mode) != HSB_LIT_SUCCESS) return fail(err);
    void *mem = nullptr;
    if (posix_memalign(&mem, 64, image.size())) return fail("out of memory");
    std::memcpy(mem, image.data(), image.size());
    *db = (hs_database_t *)mem;
    return HSB_LIT_SUCCESS;
}

extern "C" int hsb_compile_regex_rose(const char *const *expressions, const unsigned *flags, const unsigned *ids,
                                      unsigned elements, unsigned mode, unsigned engine, unsigned dfa_states,
                                      hs_database_t **db, char **error) {
    if (error) *error = nullptr;
    if (!db) return HSB_LIT_COMPILER_ERROR;
    *db = nullptr;
    auto fail = [&](const std::string &m) {
        if (error) {
            char *p = (char *)std::malloc(m.size() + 1);
            if (p) std::memcpy(p, m.c_str(), m.size() + 1);
            *error = p;
        }
        return HSB_LIT_COMPILER_ERROR;
    };
    if (!expressions || !elements) return fail("invalid parameter: no expressions");
    if (mode != hsb::MODE_BLOCK) return fail("only HS_MODE_BLOCK databases are written");
    std::vector<hsb_nfadb::Pattern> pats;
    for (unsigned i = 0; i < elements; i++) {
        if (!expressions[i]) return fail("null expression");
        pats.push_back({expressions[i], flags ? flags[i] : 0, ids ? ids[i] : 0});
    }
    std::vector<hsb::u8> image;
    std::string err;
    if (hsb_nfadb::build_rose(pats, engine, dfa_states, image, err) != HSB_LIT_SUCCESS) return fail(err);
    void *mem = nullptr;
    if (posix_memalign(&mem, 64, image.size())) return fail("out of memory");
    std::memcpy(mem, image.data(), image.size());
    *db = (hs_database_t *)mem;
    return HSB_LIT_SUCCESS;
}

extern "C" int hsb_compile_regex_multi(const char *const *expressions, const unsigned *flags, const unsigned *ids,
                                       unsigned elements, unsigned mode, hs_database_t **db, char **error) {
    return hsb_compile_regex_multi_ex(expressions, flags, ids, elements, mode, HSB_REGEX_ENGINE_LIMEX, db, error);
}
