// tests/turbo_emul.cpp -- TEST INFRASTRUCTURE: evaluates the literal front end's upload-time tables
// (vectorscan_b200/csrc/turbo_tables.h) on the host, position by position, the way k_turbo2 reads them, so that the
// soundness of the table BUILD (no literal end may be filtered out) can be checked without a GPU.
#include "turbo_tables.h"

#include <cstring>
#include <string>

using namespace hsb;

namespace {
struct Text {
    const u8 *p; u64 n;
    u32 at(long long i) const { return (i < 0 || (u64)i >= n) ? 0u : p[i]; }   // zero pads either side, as in HBM
    u32 window(long long i) const { return at(i) | at(i + 1) << 8 | at(i + 2) << 16 | at(i + 3) << 24; }
};
bool tier1_flag(const turbo::Tables &t, const Text &tx, long long pos) {
    const u32 K = t.K;
    // even positions take the A test (index = first byte of the gram), odd ones the B test (index = last byte)
    const bool a = (pos & 1) == 0;
    const long long shared_first = a ? pos - (long long)K + 2 : pos - (long long)K + 1;
    const u32 idx_byte = a ? tx.at(pos - (long long)K + 1) : tx.at(pos);
    const u32 row = K > 1 ? turbo::gram_row(tx.window(shared_first), t.gmul, t.G1) : 0u;
    const u32 word = t.gram[row];
    return ((word << (idx_byte & 31)) & 0x80000000u) != 0;
}
bool tier2_flag(const turbo::Tables &t, const Text &tx, long long pos) {
    u32 ktail = 0;
    for (u32 j = 0; j < t.K; j++) ktail |= tx.at(pos - (long long)t.K + 1 + j) << (8 * j);
    const u32 extra = tx.at(pos - (long long)t.K);
    const u32 h = turbo::tail_hash(ktail), w = t.bloom[turbo::tail_word(h, t.BL)];
    u32 r = w >> (h & 31);
    if (t.L) r |= w >> (extra & 31);
    return r & 1;
}
}  // namespace

extern "C" {
// ends[i] = index of a literal's LAST byte in the flat text.  Returns the number of ends either tier would drop
// (must be 0), or -1 when the tables cannot be built.  stats: [0] K, [1] wide form?, [2] G1, [3] tier-1 survivors,
// [4] tier-2 survivors over the whole text.
long turbo_emul_check(const void *db, const u8 *text, u64 len, const u64 *ends, u32 n_ends, u64 *stats) {
    DbView v;
    std::string why;
    if (parse_database(db, v, why) != E_SUCCESS) return -1;
    turbo::Tables t;
    if (!turbo::build(v, t)) return -1;
    const Text tx{text, len};
    long dropped = 0;
    for (u32 i = 0; i < n_ends; i++)
        if (!tier1_flag(t, tx, (long long)ends[i]) || !tier2_flag(t, tx, (long long)ends[i])) dropped++;
    u64 c1 = 0, c2 = 0;
    for (u64 p = 0; p < len; p++)
        if (tier1_flag(t, tx, (long long)p)) { c1++; if (tier2_flag(t, tx, (long long)p)) c2++; }
    if (stats) { stats[0] = t.K; stats[1] = t.wide; stats[2] = t.G1; stats[3] = c1; stats[4] = c2; }
    return dropped;
}
}
