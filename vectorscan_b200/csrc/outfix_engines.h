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
}  // namespace hsb_nfadb
