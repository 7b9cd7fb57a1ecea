// outfix_engines.h -- what litdb_writer.cpp (mixed literal + regex databases) needs from nfadb_writer.cpp
#pragma once
#include "hsb_layout.h"

#include <string>
#include <vector>

namespace hsb_nfadb {
struct Pattern { std::string expr; unsigned flags; unsigned id; };
// one outfix engine, relative to its own start: struct NFA + implementation (LimEx or McClellan)
struct EngineBlob { std::vector<hsb::u8> bytes; hsb::u32 stream_state = 0, scratch_state = 0, min_width = 0; };
// Groups the patterns into outfix engines (McClellan DFAs of <= dfa_states states, LimEx NFAs for what does not
// determinise; engine = HSB_REGEX_ENGINE_*) and lays every engine out.  program_of[i] = bytecode offset of pattern i's
// report program (the "ReportID" its engine raises).
int build_outfix_engines(const std::vector<Pattern> &pats, unsigned engine, hsb::u32 dfa_states,
                         const std::vector<hsb::u32> &program_of, std::vector<EngineBlob> &out, std::string &err);

inline bool accepts_eod(const EngineBlob &b) { return (hsb::rd32(b.bytes.data(), hsb::nfa::flags) & hsb::nfa::ACCEPTS_EOD) != 0; }

// The EOD program of a database whose outfix engines have end-of-data accepts (patterns ending in \z):
//     ENGINES_EOD {iter_offset}; END        (makeEodProgram / addEodEventProgram, src/rose/rose_build_program.cpp; run by
//     roseBlockEodExec, src/rose/block.c:163-190, after every engine was caught up to the end of the block)
// ENGINES_EOD walks the queues named by a multibit SPARSE ITERATOR that are still active and runs nfaCheckFinalState on
// each (roseEnginesEod, src/rose/program_runtime.c:1739-1789).  Iterator records {u64 mask; u32 val} as
// mmbBuildSparseIterator lays them out for the flat multibit model (<= 256 keys; src/util/multibit.h:1136-1184): up to 64
// keys -> one record whose mask holds the keys; more -> a root record (mask = 64-key blocks that hold keys, val = index
// of the first block record) followed by one record per such block (mask = its keys, val = keys in earlier blocks).
// Returns the program's offset, 0 when no engine accepts at EOD; `rose` is either writer's byte buffer.
template <class BufT>
hsb::u32 emit_engines_eod(BufT &rose, const std::vector<EngineBlob> &blobs, std::string &err) {
    using namespace hsb;
    const u32 n = (u32)blobs.size();
    std::vector<u32> keys;
    for (u32 q = 0; q < n; q++) if (accepts_eod(blobs[q])) keys.push_back(q);
    if (keys.empty()) return 0;
    if (n > 256) { err = "end-of-data patterns in a database of more than 256 engines are not written"; return ~0u; }
    rose.align(8);
    const u32 iter_off = (u32)rose.size();
    auto put = [&](u64 mask, u32 val) { const size_t o = rose.grow(16); wr64(rose.at(o), 0, mask); wr32(rose.at(o), 8, val); };
    if (n <= 64) {
        u64 mask = 0;
        for (u32 k : keys) mask |= 1ull << k;
        put(mask, 0);
    } else {
        u64 root = 0, block_mask[4] = {0, 0, 0, 0};
        for (u32 k : keys) { root |= 1ull << (k / 64); block_mask[k / 64] |= 1ull << (k % 64); }
        put(root, 1);
        u32 before = 0;
        for (u32 b = 0; b < 4; b++)
            if (block_mask[b]) { put(block_mask[b], before); before += (u32)__builtin_popcountll(block_mask[b]); }
    }
    rose.align(op::ALIGN);
    const u32 prog = (u32)rose.size();
    const size_t o = rose.grow(8 + op::SIZE_1);
    wr8(rose.at(o), 0, op::ENGINES_EOD);
    wr32(rose.at(o), 4, iter_off);
    wr8(rose.at(o + 8), 0, op::END);
    return prog;
}
}  // namespace hsb_nfadb
