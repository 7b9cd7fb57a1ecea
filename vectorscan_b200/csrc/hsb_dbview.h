// hsb_dbview.h -- host-side reader of a block-mode hs_database_t: header checks with the reference's
// error codes (src/runtime.c:316-340, src/database.c:330-364), location of the literal matcher and its
// confirm structures, and the ENGINE / OPCODE CENSUS that refuses (HS_ARCH_ERROR, never a fallback) any
// database using something the device path does not implement.
#pragma once
#include "hsb_layout.h"

#include <algorithm>
#include <set>
#include <string>
#include <vector>

namespace hsb {

constexpr int E_SUCCESS = 0, E_INVALID = -1, E_NOMEM = -2, E_DB_VERSION = -5, E_DB_PLATFORM = -6,
              E_DB_MODE = -7, E_BAD_ALIGN = -8, E_ARCH = -11, E_UNKNOWN = -13;

enum EngineKind : u32 { ENG_FDR = 0, ENG_TEDDY = 1, ENG_NOODLE = 2, ENG_LIMEX = 3, ENG_MCCLELLAN = 4, ENG_SHENG = 5,
                        ENG_MULTI = 6 /* full-Rose database of outfix engines: DbView::engines */ };

struct LitRec {  // one LitInfo, flattened
    u64 v, msk, groups;
    u32 program;  // Rose program offset (LitInfo.id)
    u8 size, flags, bucket;
};

struct DbView {
    const u8 *bc = nullptr;  // bytecode (RoseEngine) base inside the caller's database
    u32 bc_len = 0, crc = 0;
    u32 min_width = 0, min_width_excl = 0, max_bi_width = 0;
    u32 ekey_count = 0, dkey_count = 0;
    u64 groups = 0;  // initialGroups & floating_group_mask
    u32 hwlm_off = 0, eng_off = 0;
    EngineKind kind = ENG_FDR;
    u32 engine_id = 0, n_buckets = 8, max_string_len = 0, n_strings = 0;
    u32 conf_base = 0;  // absolute offset of u32 confBase[n_buckets]
    // FDR
    u32 fdr_tab = 0, fdr_domain = 0, fdr_stride = 0;
    u64 fdr_start = 0;
    // Teddy
    u32 teddy_masks = 0, teddy_mask_off = 0, teddy_fat = 0;
    // noodle
    u64 nood_msk = 0, nood_cmp = 0;
    u32 nood_len = 0, nood_id = 0;
    std::vector<LitRec> lits;  // every literal of every bucket (empty for noodle)
    std::vector<u32> programs; // distinct program offsets reachable from the literal matcher
    u32 max_lit_len = 0;       // longest CHECK_*_LIT length seen (bounds look-behind)
    // single-outfix LimEx NFA (kind == ENG_LIMEX): offsets are absolute bytecode offsets
    u32 nfa_off = 0, limex_off = 0, limex_type = 0, nfa_flags = 0, nfa_len = 0;
    // single-outfix McClellan DFA (kind == ENG_MCCLELLAN), absolute bytecode offsets
    struct Mcc {
        u32 m_off = 0, tab_off = 0, tab_bytes = 0, aux_off = 0, sherman_off = 0;
        u32 wide16 = 0, as = 0, states = 0, start = 0, accept_limit_8 = 0, sherman_limit = 0, single = 0, arb_report = 0;
        u32 remap_off = 0;                  // absolute offset of remap[256]
        u32 sheng_end = 0, masks_off = 0, mask_stride = 16;   // McSheng: first state outside the Sheng region, absolute offset of sheng_masks, bytes per mask (64: mcsheng64)
        // wide states (16-bit McClellan): ids >= wide_limit are heads of symbol chains stored at wide_off (absolute);
        // wide_rows = sum of the chains' widths (every chain position becomes a table row on the device)
        u32 has_wide = 0, wide_limit = 0, wide_off = 0, wide_number = 0, wide_rows = 0, n_alpha = 0;
    } mcc;
    // single-outfix Sheng DFA (kind == ENG_SHENG), absolute bytecode offsets
    struct Sheng { u32 sh_off = 0, aux_off = 0, n_states = 0, anchored = 0, flags = 0, report = 0, nfa_type = 17; } sheng;
    // full-Rose database of outfix engines (kind == ENG_MULTI): one view per queue, each of kind LIMEX / MCCLELLAN
    std::vector<DbView> engines;
    u32 nfa_min_width = 0;  // NFA.minWidth of an outfix engine: blocks shorter than this do not start it
    // HS_MODE_VECTORED (src/runtime.c:1106-1174): the buffers of one hs_scan_vector call are consecutive blocks of one
    // stream -- offsets run on, a literal may straddle blocks, exhaustion lasts to the end of the call
    bool vectored = false;
    // HS_MODE_STREAM (hs_open_stream / hs_scan_stream / hs_close_stream, src/runtime.c:543-1060): pure-literal databases
    // only.  A block is scanned behind the stream's last historyRequired bytes (kept on the host with the stream); what
    // must survive from call to call beside that is the exhaustion of single-match reports, which the host keeps per
    // stream by exhaustion key: stream_ekey_of maps an external report id to its key (~0u: the report never exhausts).
    bool streaming = false;
    u32 history_required = 0;
    std::vector<std::pair<u32, u32>> stream_ekey_of;
    bool stream_any_ekey = false;
};

inline bool in_range(const DbView &d, size_t off, size_t n) { return off <= d.bc_len && n <= d.bc_len - off; }

// Padded size of an instruction the device interpreter implements (0 = not implemented) and the offset of its
// fail_jump operand (0 = the instruction cannot fail).
inline size_t op_size(u8 code, size_t &fail_jump_at) {
    fail_jump_at = 0;
    switch (code) {
    case op::END: case op::CATCH_UP: case op::CLEAR_WORK_DONE: return op::SIZE_1;
    case op::CHECK_GROUPS: return op::check_groups::SIZE;
    case op::CHECK_MASK: fail_jump_at = op::check_mask::fail_jump; return op::check_mask::SIZE;
    case op::CHECK_MASK_32: fail_jump_at = op::check_mask32::fail_jump; return op::check_mask32::SIZE;
    case op::CHECK_BYTE: fail_jump_at = op::check_byte::fail_jump; return op::check_byte::SIZE;
    case op::CHECK_SHUFTI_16x8: fail_jump_at = op::check_shufti_16x8::fail_jump; return op::check_shufti_16x8::SIZE;
    case op::CHECK_SHUFTI_32x8: fail_jump_at = op::check_shufti_32x8::fail_jump; return op::check_shufti_32x8::SIZE;
    case op::CHECK_SHUFTI_16x16: fail_jump_at = op::check_shufti_16x16::fail_jump; return op::check_shufti_16x16::SIZE;
    case op::CHECK_SHUFTI_32x16: fail_jump_at = op::check_shufti_32x16::fail_jump; return op::check_shufti_32x16::SIZE;
    case op::CHECK_MASK_64: fail_jump_at = op::check_mask64::fail_jump; return op::check_mask64::SIZE;
    case op::CHECK_SHUFTI_64x8: fail_jump_at = op::check_shufti_64x8::fail_jump; return op::check_shufti_64x8::SIZE;
    case op::CHECK_SHUFTI_64x16: fail_jump_at = op::check_shufti_64x16::fail_jump; return op::check_shufti_64x16::SIZE;
    case op::DEDUPE: fail_jump_at = op::dedupe::fail_jump; return op::dedupe::SIZE;
    case op::REPORT: case op::FINAL_REPORT: return op::report::SIZE;
    case op::REPORT_EXHAUST: return op::report_exhaust::SIZE;
    case op::DEDUPE_AND_REPORT: fail_jump_at = op::dedupe_and_report::fail_jump; return op::dedupe_and_report::SIZE;
    case op::CHECK_EXHAUSTED: fail_jump_at = op::check_exhausted::fail_jump; return op::check_exhausted::SIZE;
    case op::SQUASH_GROUPS: return op::squash_groups::SIZE;
    case op::SET_EXHAUST: return op::set_exhaust::SIZE;
    case op::INCLUDED_JUMP: return op::included_jump::SIZE;
    case op::CHECK_LONG_LIT: case op::CHECK_LONG_LIT_NOCASE: case op::CHECK_MED_LIT: case op::CHECK_MED_LIT_NOCASE:
        fail_jump_at = op::check_lit::fail_jump; return op::check_lit::SIZE;
    default: return 0;
    }
}

// Walk one Rose program to its END; this is the validation gate for caller-supplied blobs, so every instruction must lie
// inside the bytecode with its full operand block, every fail_jump must land on an instruction of this same program at
// or before its END (the device interpreter follows them unchecked), and a DEDUPE must be followed by one of the three
// reports the interpreter fuses it with.  INCLUDED_JUMP children are programs of other literals, walked on their own.
inline int census_program(const DbView &d, u32 prog, std::string &why, u32 &max_lit_len, u64 *hist = nullptr,
                          std::vector<std::pair<u32, u32>> *reports = nullptr) {  // reports: (external id, ekey or ~0u) of every reporting instruction
    size_t pc = prog;
    std::vector<size_t> starts, targets;
    for (int guard = 0; guard < 4096; guard++) {
        if (!in_range(d, pc, 8) || (pc & 7)) { why = "program runs past the bytecode"; return E_INVALID; }
        const u8 code = rd8(d.bc, pc);
        size_t fj = 0;
        const size_t sz = op_size(code, fj);
        if (!sz) {
            why = "Rose opcode " + std::to_string((int)code) + " is not implemented on the device path";
            return E_ARCH;
        }
        if (!in_range(d, pc, sz)) { why = "instruction operands run past the bytecode"; return E_INVALID; }
        if (hist) hist[code & 127]++;
        if (reports) {
            if (code == op::REPORT || code == op::FINAL_REPORT) reports->push_back({rd32(d.bc, pc + op::report::onmatch), ~0u});
            if (code == op::DEDUPE_AND_REPORT) reports->push_back({rd32(d.bc, pc + op::dedupe_and_report::onmatch), ~0u});
            if (code == op::REPORT_EXHAUST)
                reports->push_back({rd32(d.bc, pc + op::report_exhaust::onmatch), rd32(d.bc, pc + op::report_exhaust::ekey)});
        }
        starts.push_back(pc);
        if (fj) {
            const u32 j = rd32(d.bc, pc + fj);
            if (j < sz || (j & 7)) { why = "fail_jump does not move forward to an instruction boundary"; return E_INVALID; }
            targets.push_back(pc + j);
        }
        if (code == op::CHECK_LONG_LIT || code == op::CHECK_LONG_LIT_NOCASE || code == op::CHECK_MED_LIT ||
            code == op::CHECK_MED_LIT_NOCASE) {
            const u32 lo = rd32(d.bc, pc + op::check_lit::lit_offset), ll = rd32(d.bc, pc + op::check_lit::lit_length);
            if (!in_range(d, lo, ll)) { why = "literal text outside the bytecode"; return E_INVALID; }
            if (ll > max_lit_len) max_lit_len = ll;
        }
        if (code == op::DEDUPE) {
            const u8 nxt = in_range(d, pc + sz, 8) ? rd8(d.bc, pc + sz) : (u8)op::END;
            if (nxt != op::REPORT && nxt != op::REPORT_EXHAUST && nxt != op::SET_EXHAUST) {
                why = "DEDUPE followed by opcode " + std::to_string((int)nxt) + " is not implemented on the device path";
                return E_ARCH;
            }
        }
        if (code == op::END || code == op::FINAL_REPORT) {  // FINAL_REPORT always terminates the program
            // a FINAL_REPORT in mid-program is followed by the blocks its predecessors' fail_jumps skip to: keep walking
            // while some jump still points beyond this instruction
            bool beyond = false;
            for (size_t t : targets) if (t > pc) beyond = true;
            if (code == op::END || !beyond) {
                for (size_t t : targets) {
                    bool found = false;
                    for (size_t st : starts) if (st == t) found = true;
                    if (!found) { why = "fail_jump target is not an instruction of its program"; return E_INVALID; }
                }
                return E_SUCCESS;
            }
        }
        pc += sz;
    }
    why = "program too long";
    return E_INVALID;
}


// Single-outfix McClellan DFA (mcclellan_internal.h:77-101).  Accepted: 8- and 16-bit tables, Sherman states,
// per-state report lists, EOD accepts, wide states of 16-bit engines.  Accel structures are skip-ahead only.
inline int parse_mcclellan(DbView &d, std::string &why) {
    namespace mc = mcclellan;
    namespace ms = mcsheng;
    const u8 *r = d.bc;
    const u8 *n = r + d.nfa_off;
    DbView::Mcc &M = d.mcc;
    // McSheng (types 19 / 20) is McClellan with another header and a Sheng region in front (hsb_layout.h)
    const u8 type = rd8(n, nfa::type);
    const bool hybrid64 = type == nfa::MCSHENG64_8 || type == nfa::MCSHENG64_16;   // struct mcsheng64: 64-byte masks
    const bool hybrid = type == nfa::MCSHENG_8 || type == nfa::MCSHENG_16 || hybrid64;
    const size_t HDR = hybrid64 ? ms::SIZE64 : hybrid ? ms::SIZE : mc::SIZE;
    auto F = [&](size_t plain, size_t hyb) { return hybrid ? hyb : plain; };
    if (d.nfa_len < nfa::SIZE + HDR) { why = "McClellan truncated"; return E_INVALID; }
    M.m_off = d.nfa_off + (u32)nfa::SIZE;
    const u8 *m = r + M.m_off;
    auto n_range = [&](size_t off, size_t len) { return off <= d.nfa_len && len <= d.nfa_len - off; };  // relative to NFA
    M.wide16 = type == nfa::MCCLELLAN_16 || type == nfa::MCSHENG_16 || type == nfa::MCSHENG64_16;
    M.has_wide = !hybrid && rd8(m, mc::has_wide) ? 1 : 0;
    if (M.has_wide && !M.wide16) { why = "wide states in an 8-bit McClellan engine are not supported yet"; return E_ARCH; }
    M.as = rd8(m, F(mc::alphaShift, ms::alphaShift));
    M.states = rd16(m, mc::state_count);
    M.start = rd16(m, mc::start_anchored);
    M.accept_limit_8 = rd16(m, F(mc::accept_limit_8, ms::accept_limit_8));
    M.sherman_limit = M.wide16 ? rd16(m, F(mc::sherman_limit, ms::sherman_limit)) : M.states;
    M.single = (rd8(m, F(mc::flags, ms::flags)) & mc::FLAG_SINGLE) ? 1 : 0;
    M.arb_report = rd32(m, F(mc::arb_report, ms::arb_report));
    M.remap_off = M.m_off + (u32)F(mc::remap, ms::remap);
    M.sheng_end = hybrid ? rd16(m, ms::sheng_end) : 0;
    M.masks_off = hybrid ? M.m_off + (u32)(hybrid64 ? ms::sheng_succ_masks64 : ms::sheng_masks) : 0;
    M.mask_stride = hybrid64 ? 64 : 16;
    if (M.as > 8 || !M.states || M.start >= M.states || M.sherman_limit > M.states) { why = "McClellan header inconsistent"; return E_INVALID; }
    // states with an mstate_aux record and a Sherman slot end at wide_limit; ids a table may name end behind the wide heads
    u32 n_aux = M.states, ref_limit = M.states;
    if (M.has_wide) {
        M.wide_limit = rd16(m, mc::wide_limit);
        const u32 w_rel = rd32(m, mc::wide_offset);
        if (M.wide_limit < M.sherman_limit || M.wide_limit >= M.states || (w_rel & 1) || !n_range(w_rel, 8)) { why = "McClellan wide region out of range"; return E_INVALID; }
        M.wide_off = d.nfa_off + w_rel;
        M.wide_number = rd16(r, M.wide_off + 2);
        if (rd8(r, M.wide_off) != 2 || !M.wide_number || M.wide_limit + M.wide_number > M.states || !n_range(w_rel, 4 + 4 * (size_t)M.wide_number)) {
            why = "McClellan wide region malformed"; return E_INVALID;
        }
        n_aux = M.wide_limit;
        ref_limit = M.wide_limit + M.wide_number;
    } else {
        M.wide_limit = M.states;   // (no wide states: every id below `states` has its aux record)
    }
    for (u32 c = 0; c < 256; c++) M.n_alpha = std::max<u32>(M.n_alpha, rd8(r, M.remap_off + c) + 1u);
    if (M.wide16 && M.states - 1 > mc::STATE_MASK) { why = "McClellan state count out of range"; return E_INVALID; }
    if (!M.wide16 && (M.states > 256 || M.accept_limit_8 > M.states)) { why = "McClellan8 limits out of range"; return E_INVALID; }
    if (hybrid && (M.sheng_end < 2 || M.sheng_end > (hybrid64 ? ms::MAX_SHENG64_STATES : ms::MAX_SHENG_STATES) + 1 || M.sheng_end > M.sherman_limit)) { why = "McSheng region out of range"; return E_INVALID; }
    const u32 rows = (M.wide16 ? M.sherman_limit : M.states) - M.sheng_end;   // the table has no rows for the Sheng region
    if (M.has_wide && M.start >= ref_limit) { why = "McClellan start state out of range"; return E_INVALID; }
    M.tab_off = M.m_off + (u32)HDR;
    M.tab_bytes = (rows << M.as) * (M.wide16 ? 2u : 1u);
    if (!n_range(nfa::SIZE + HDR, M.tab_bytes)) { why = "McClellan successor table truncated"; return E_INVALID; }
    for (u32 c = 0; c < 256; c++)
        if (rd8(r, M.remap_off + c) >= (1u << M.as)) { why = "McClellan remap out of range"; return E_INVALID; }
    const u32 aux_rel = rd32(m, mc::aux_offset);
    if (!n_range(aux_rel, (size_t)n_aux * mc::aux::SIZE)) { why = "McClellan aux array truncated"; return E_INVALID; }
    M.aux_off = d.nfa_off + aux_rel;
    // every table entry names a state
    for (u32 i = 0; i < (rows << M.as); i++) {
        const u32 e = M.wide16 ? (rd16(r, M.tab_off + 2 * (size_t)i) & mc::STATE_MASK) : rd8(r, M.tab_off + i);
        if (e >= ref_limit) { why = "McClellan successor out of range"; return E_INVALID; }
    }
    if (hybrid) {
        // mask byte: < sheng_end - 1 a region state (id - 1), == sheng_end - 1 dead, above it a state id outside the region
        for (u32 c = 0; c < 256; c++)
            for (u32 sid = 0; sid + 1 < M.sheng_end; sid++)
                if (rd8(r, M.masks_off + M.mask_stride * c + sid) >= M.states) { why = "McSheng mask successor out of range"; return E_INVALID; }
        for (u32 st = 1; st < M.sheng_end; st++)
            if (rd32(r, M.aux_off + st * mc::aux::SIZE + mc::aux::accept)) { why = "McSheng region state with reports"; return E_INVALID; }
    }
    if (M.wide16 && M.sherman_limit < n_aux) {
        const u32 sh_rel = rd32(m, mc::sherman_offset);
        const size_t nsh = n_aux - M.sherman_limit;
        if ((sh_rel & 15) || !n_range(sh_rel, nsh * mc::SHERMAN_FIXED_SIZE)) { why = "McClellan Sherman region truncated"; return E_INVALID; }
        M.sherman_off = d.nfa_off + sh_rel;
        for (size_t k = 0; k < nsh; k++) {
            const u8 *sh = r + M.sherman_off + k * mc::SHERMAN_FIXED_SIZE;
            const u32 len = sh[mc::SHERMAN_LEN_OFFSET];
            const u32 daddy = rd16(sh, mc::SHERMAN_DADDY_OFFSET);
            if (sh[0] != mc::SHERMAN_STATE || len > mc::MAX_SHERMAN_LIST_LEN || daddy >= M.sherman_limit || daddy < M.sheng_end) {
                why = "McClellan Sherman state malformed"; return E_INVALID;
            }
            for (u32 q = 0; q < len; q++)
                if ((rd16(sh, mc::SHERMAN_CHARS_OFFSET + len + 2 * q) & mc::STATE_MASK) >= ref_limit) { why = "McClellan Sherman successor out of range"; return E_INVALID; }
        }
    }
    if (M.has_wide) {
        // every entry: u16 width; u8 symbols[width]; pad to 2; u16 next; u16 failure[n_alpha]   (mcclellan_internal.h:53-61)
        for (u32 k = 0; k < M.wide_number; k++) {
            const u32 eo = rd32(r, M.wide_off + 4 + 4 * k);
            const size_t e_rel = (size_t)(M.wide_off - d.nfa_off) + eo;
            if ((eo & 1) || !n_range(e_rel, 2)) { why = "McClellan wide entry out of range"; return E_INVALID; }
            const u32 width = rd16(r, M.wide_off + eo);
            const size_t tr = 2 + (((size_t)width + 1) & ~(size_t)1);
            if (!width || !n_range(e_rel, tr + 2 * (1 + (size_t)M.n_alpha))) { why = "McClellan wide entry truncated"; return E_INVALID; }
            for (u32 q = 0; q < width; q++)
                if (rd8(r, M.wide_off + eo + 2 + q) >= (1u << M.as)) { why = "McClellan wide symbol out of range"; return E_INVALID; }
            for (u32 q = 0; q < 1 + M.n_alpha; q++)
                if ((rd16(r, M.wide_off + eo + tr + 2 * q) & mc::STATE_MASK) >= ref_limit) { why = "McClellan wide successor out of range"; return E_INVALID; }
            M.wide_rows += width;
        }
        if (M.wide_limit + M.wide_rows - 1 > mc::STATE_MASK) { why = "McClellan wide states: more than 16384 table rows"; return E_ARCH; }
    }
    std::set<u32> progs;
    if (M.single) progs.insert(M.arb_report);
    for (u32 st = 0; st < n_aux; st++) {
        for (size_t f : {mc::aux::accept, mc::aux::accept_eod}) {
            const u32 rl = rd32(r, M.aux_off + st * mc::aux::SIZE + f);
            if (!rl) continue;
            if ((rl & 3) || !n_range(rl, 4)) { why = "McClellan report list outside the engine"; return E_INVALID; }
            const u32 cnt = rd32(n, rl);
            if (cnt > 65536 || !n_range((size_t)rl + 4, 4 * (size_t)cnt)) { why = "McClellan report list truncated"; return E_INVALID; }
            for (u32 k = 0; k < cnt; k++) progs.insert(rd32(n, rl + 4 + 4 * k));
        }
    }
    d.kind = ENG_MCCLELLAN;
    d.max_lit_len = 8;
    for (u32 p : progs) {
        if (p < rose::SIZE || !in_range(d, p, 8)) { why = "report program offset outside bytecode"; return E_INVALID; }
        int rv = census_program(d, p, why, d.max_lit_len);
        if (rv != E_SUCCESS) return rv;
        d.programs.push_back(p);
    }
    return E_SUCCESS;
}


// Single-outfix Sheng DFA (sheng_internal.h:56-79): <= 16 states, 256 shuffle masks.  Accel structures are skip-ahead
// only.  A state carrying the DEAD flag must be a trap without reports (that is what the flag means; the device path
// stops there).
inline int parse_sheng(DbView &d, std::string &why) {
    const sheng::Shape sh = sheng::shape(d.limex_type);
    const u8 *r = d.bc;
    const u8 *n = r + d.nfa_off;
    if (d.nfa_len < nfa::SIZE + sh.SIZE) { why = "Sheng truncated"; return E_INVALID; }
    DbView::Sheng &S = d.sheng;
    S.sh_off = d.nfa_off + (u32)nfa::SIZE;
    const u8 *m = r + S.sh_off;
    auto n_range = [&](size_t off, size_t len) { return off <= d.nfa_len && len <= d.nfa_len - off; };
    S.n_states = rd8(m, sh.n_states);
    S.anchored = rd8(m, sh.anchored);
    S.flags = rd8(m, sh.flags);
    S.report = rd32(m, sh.report);
    if (!S.n_states || S.n_states > sh.max_states || (S.anchored & sh.mask) >= S.n_states) { why = "Sheng header inconsistent"; return E_INVALID; }
    const u32 aux_rel = rd32(m, sh.aux_offset);
    if (!n_range(aux_rel, (size_t)S.n_states * sheng::aux::SIZE)) { why = "Sheng aux array truncated"; return E_INVALID; }
    S.aux_off = d.nfa_off + aux_rel;
    std::set<u32> progs;
    if (S.flags & sheng::FLAG_SINGLE_REPORT) progs.insert(S.report);
    bool has_reports[64] = {false};
    for (u32 st = 0; st < S.n_states; st++)
        for (size_t f : {sheng::aux::accept, sheng::aux::accept_eod}) {
            const u32 rl = rd32(r, S.aux_off + st * sheng::aux::SIZE + f);
            if (!rl) continue;
            has_reports[st] = true;
            if ((rl & 3) || !n_range(rl, 4)) { why = "Sheng report list outside the engine"; return E_INVALID; }
            const u32 cnt = rd32(n, rl);
            if (cnt > 65536 || !n_range((size_t)rl + 4, 4 * (size_t)cnt)) { why = "Sheng report list truncated"; return E_INVALID; }
            for (u32 k = 0; k < cnt; k++) progs.insert(rd32(n, rl + 4 + 4 * k));
        }
    for (u32 c = 0; c < 256; c++)
        for (u32 st = 0; st < S.n_states; st++) {
            const u8 e = rd8(m, (size_t)sh.stride * c + st);
            const u32 t = e & sh.mask;
            if (t >= S.n_states) { why = "Sheng successor out of range"; return E_INVALID; }
            if (((e & sh.accept) != 0) != (rd32(r, S.aux_off + t * sheng::aux::SIZE + sheng::aux::accept) != 0)) {
                why = "Sheng accept flag disagrees with the state's report list"; return E_INVALID;
            }
            if (e & sh.dead) {
                if (has_reports[t]) { why = "Sheng dead state with reports"; return E_INVALID; }
                for (u32 c2 = 0; c2 < 256; c2++)
                    if ((rd8(m, (size_t)sh.stride * c2 + t) & sh.mask) != t) { why = "Sheng dead state is not a trap"; return E_INVALID; }
            }
        }
    S.nfa_type = d.limex_type;
    d.kind = ENG_SHENG;
    d.max_lit_len = 8;
    for (u32 p : progs) {
        if (p < rose::SIZE || !in_range(d, p, 8)) { why = "report program offset outside bytecode"; return E_INVALID; }
        int rv = census_program(d, p, why, d.max_lit_len);
        if (rv != E_SUCCESS) return rv;
        d.programs.push_back(p);
    }
    return E_SUCCESS;
}

// Single-outfix database (src/runtime.c:246-279): queue 0 holds the only engine.  Accepted when that engine is a
// LimEx NFA without bounded repeats and without reverse acceleration; every report list and accept entry must
// lead to a Rose report program the device interpreter implements.
inline int parse_outfix_at(DbView &d, u32 info_off, std::string &why);

inline int parse_single_outfix(DbView &d, std::string &why) {
    const u8 *r = d.bc;
    if (rd32(r, rose::fmatcherOffset) || rd32(r, rose::amatcherOffset) || rd32(r, rose::ematcherOffset)) {
        why = "single-outfix database with a literal matcher"; return E_INVALID;
    }
    if (rd32(r, rose::outfixEndQueue) != 1 || rd32(r, rose::queueCount) != 1) { why = "single-outfix database with more than one queue"; return E_INVALID; }
    return parse_outfix_at(d, rd32(r, rose::nfaInfoOffset), why);
}

// Full-Rose database whose only engines are outfixes (src/rose/block.c:345-422 with no literal matcher: every outfix is
// started by blockInitSufPQ and caught up to the end of the block, src/rose/catchup.c:560-672).  Accepted when there is
// nothing else for Rose to orchestrate: no anchored / floating / EOD / delay-rebuild / small-block matcher, no roles, no
// leftfixes, no MPV, no EOD program, no delayed or anchored literals.  Each queue's engine is censused like a sole outfix.
inline int full_rose_gates(const DbView &d, std::string &why) {
    const u8 *r = d.bc;
    if (rd32(r, rose::amatcherOffset) || rd32(r, rose::ematcherOffset) || rd32(r, rose::drmatcherOffset) ||
        rd32(r, rose::sbmatcherOffset) || rd32(r, rose::longLitTableOffset)) {
        why = "full-Rose databases with anchored / EOD / delay-rebuild / small-block matchers or long literals are not supported yet";
        return E_ARCH;
    }
    const u32 nq = rd32(r, rose::queueCount);
    if (rd32(r, rose::rolesWithStateCount) || rd32(r, rose::activeLeftCount) || rd32(r, rose::roseCount) ||
        rd32(r, rose::delay_count) ||
        rd32(r, rose::anchored_count) || rd32(r, rose::initMpvNfa) != 0xffffffffu || rd32(r, rose::outfixBeginQueue) ||
        rd32(r, rose::outfixEndQueue) != nq || rd32(r, rose::leftfixBeginQueue) != nq || rd32(r, rose::somLocationCount)) {
        why = "full-Rose database with roles, leftfixes, MPV or delayed literals is not supported yet";
        return E_ARCH;
    }
    // EOD program: only "ENGINES_EOD; END" -- the end-of-data accepts of the outfix engines (roseEnginesEod,
    // src/rose/program_runtime.c:1739-1789), which every engine kernel raises itself at the end of a buffer
    const u32 eodp = rd32(r, rose::eodProgramOffset);
    if (eodp && rd8(r, rose::requiresEodCheck)) {
        if (!in_range(d, eodp, 16) || (eodp & 7) || rd8(r, eodp) != op::ENGINES_EOD || rd8(r, eodp + 8) != op::END) {
            why = "EOD program other than ENGINES_EOD; END is not supported yet";
            return E_ARCH;
        }
    }
    if (rd32(r, rose::fmatcherOffset) &&
        (rd8(r, rose::noFloatingRoots) || rd32(r, rose::floatingMinDistance) || rd32(r, rose::floatingDistance) != rose::BOUND_INF ||
         rd32(r, rose::fmatcherMaxBiAnchoredWidth) != rose::BOUND_INF)) {
        why = "floating matcher with a distance window is not supported yet";
        return E_ARCH;
    }
    if (nq > 4096) { why = "outfix queue count out of range"; return E_INVALID; }
    return E_SUCCESS;
}

// The outfix engines of a full-Rose database (src/rose/block.c:345-422: every outfix is started by blockInitSufPQ and
// caught up to the end of the block, src/rose/catchup.c:560-672), one view per queue, each censused like a sole outfix.
inline int parse_outfix_queues(DbView &d, std::string &why) {
    const u8 *r = d.bc;
    const u32 nq = rd32(r, rose::queueCount);
    const u32 info_off = rd32(r, rose::nfaInfoOffset);
    if (!in_range(d, info_off, (size_t)nfainfo::SIZE * nq)) { why = "NfaInfo array outside bytecode"; return E_INVALID; }
    std::set<u32> progs(d.programs.begin(), d.programs.end());
    std::vector<u32> eod_queues;
    for (u32 qi = 0; qi < nq; qi++) {
        DbView e;
        e.bc = d.bc; e.bc_len = d.bc_len; e.crc = d.crc;
        e.min_width = d.min_width; e.min_width_excl = d.min_width_excl; e.max_bi_width = d.max_bi_width;
        e.ekey_count = d.ekey_count; e.dkey_count = d.dkey_count; e.groups = d.groups;
        const int rv = parse_outfix_at(e, info_off + (u32)nfainfo::SIZE * qi, why);
        if (rv != E_SUCCESS) { why = "queue " + std::to_string(qi) + ": " + why; return rv; }
        if (e.nfa_flags & nfa::ACCEPTS_EOD) eod_queues.push_back(qi);
        for (u32 p : e.programs) progs.insert(p);
        d.max_lit_len = std::max(d.max_lit_len, e.max_lit_len);
        d.engines.push_back(std::move(e));
    }
    d.programs.assign(progs.begin(), progs.end());
    // The engines' end-of-data accepts fire (in the reference) only through ENGINES_EOD and only for the queues its sparse
    // iterator names: it must name exactly the engines that accept at EOD (flat multibit model, src/util/multibit.h:1136-1184)
    if (!eod_queues.empty()) {
        const u32 eodp = rd32(r, rose::eodProgramOffset);
        if (!eodp || !rd8(r, rose::requiresEodCheck)) { why = "engines with end-of-data accepts but no EOD program"; return E_ARCH; }
        if (nq > 256) { why = "ENGINES_EOD over more than 256 queues is not supported yet"; return E_ARCH; }
        const u32 it = rd32(r, eodp + 4);
        if ((it & 7) || !in_range(d, it, 16 * 5)) { why = "sparse iterator outside bytecode"; return E_INVALID; }
        std::vector<u32> named;
        if (nq <= 64) {
            const u64 m = rd64(r, it);
            for (u32 k = 0; k < 64; k++) if ((m >> k) & 1) named.push_back(k);
        } else {
            const u64 root = rd64(r, it);
            u32 rec = rd32(r, it + 8);
            for (u32 b = 0; b < 4; b++) {
                if (!((root >> b) & 1)) continue;
                const u64 m = rd64(r, it + 16 * (size_t)rec);
                for (u32 k = 0; k < 64; k++) if ((m >> k) & 1) named.push_back(64 * b + k);
                rec++;
            }
        }
        if (named != eod_queues) { why = "ENGINES_EOD iterator does not name the engines that accept at end of data"; return E_ARCH; }
    }
    return E_SUCCESS;
}

// Full-Rose database whose only engines are outfixes: nothing else for Rose to orchestrate.
inline int parse_outfix_rose(DbView &d, std::string &why) {
    if (!rd32(d.bc, rose::queueCount)) { why = "full-Rose database with neither a literal matcher nor engines"; return E_INVALID; }
    const int rv = parse_outfix_queues(d, why);
    if (rv != E_SUCCESS) return rv;
    d.kind = ENG_MULTI;
    return E_SUCCESS;
}

inline int parse_outfix_at(DbView &d, u32 info_off, std::string &why) {
    const u8 *r = d.bc;
    if (!in_range(d, info_off, nfainfo::SIZE)) { why = "NfaInfo outside bytecode"; return E_INVALID; }
    d.nfa_off = rd32(r, info_off + nfainfo::nfaOffset);
    if ((d.nfa_off & 63) || !in_range(d, d.nfa_off, nfa::SIZE)) { why = "NFA header outside bytecode or misaligned"; return E_INVALID; }
    const u8 *n = r + d.nfa_off;
    d.nfa_len = rd32(n, nfa::length);
    d.nfa_flags = rd32(n, nfa::flags);
    d.nfa_min_width = rd32(n, nfa::minWidth);
    d.limex_type = rd8(n, nfa::type);
    if (!in_range(d, d.nfa_off, d.nfa_len)) { why = "NFA runs past the bytecode"; return E_INVALID; }
    if (rd8(n, nfa::rAccelType)) { why = "reverse acceleration is not supported"; return E_ARCH; }
    if (d.limex_type == nfa::MCCLELLAN_8 || d.limex_type == nfa::MCCLELLAN_16 || d.limex_type == nfa::MCSHENG_8 ||
        d.limex_type == nfa::MCSHENG_16 || d.limex_type == nfa::MCSHENG64_8 || d.limex_type == nfa::MCSHENG64_16)
        return parse_mcclellan(d, why);
    if (d.limex_type == nfa::SHENG || d.limex_type == nfa::SHENG_32 || d.limex_type == nfa::SHENG_64) return parse_sheng(d, why);
    if (d.limex_type > nfa::LIMEX_512) { why = "outfix engine type " + std::to_string(d.limex_type) + " (not LimEx, McClellan, McSheng or Sheng) is not supported yet"; return E_ARCH; }
    const limex::Layout L = limex::layout(d.limex_type);
    if (d.nfa_len < nfa::SIZE + L.SIZE) { why = "LimEx truncated"; return E_INVALID; }
    d.limex_off = d.nfa_off + (u32)nfa::SIZE;
    d.kind = ENG_LIMEX;
    const u8 *lx = r + d.limex_off;
    const u32 lx_len = d.nfa_len - (u32)nfa::SIZE;
    auto lx_range = [&](size_t off, size_t len) { return off <= lx_len && len <= lx_len - off; };
    if (rd32(lx, limex::repeatCount)) { why = "LimEx bounded repeats are not supported yet"; return E_ARCH; }
    if (rd32(lx, L.shiftCount) > limex::MAX_SHIFT_COUNT) { why = "LimEx shift count out of range"; return E_INVALID; }
    for (u32 k = 0; k < rd32(lx, L.shiftCount); k++)
        if (rd8(lx, L.shiftAmount + k) > limex::MAX_SHIFT_AMOUNT) { why = "LimEx shift amount out of range"; return E_INVALID; }
    if (!lx_range(L.SIZE, (size_t)rd32(lx, limex::reachSize) * L.bytes) || rd32(lx, limex::reachSize) > 256) { why = "LimEx reach table truncated"; return E_INVALID; }
    for (u32 c = 0; c < 256; c++)
        if (rd8(lx, limex::reachMap + c) >= rd32(lx, limex::reachSize)) { why = "LimEx reach map out of range"; return E_INVALID; }
    std::set<u32> progs;
    auto add_list = [&](u32 off) -> bool {  // ReportID list relative to the LimEx, MO_INVALID_IDX-terminated
        for (u32 k = 0;; k++) {
            if (k > 65536 || !lx_range((size_t)off + 4 * (size_t)k, 4)) return false;
            const u32 id = rd32(lx, (size_t)off + 4 * (size_t)k);
            if (id == 0xffffffffu) return true;
            progs.insert(id);
        }
    };
    // exception table: one entry per bit of exceptionMask
    u32 n_exc = 0;
    for (u32 b = 0; b < L.bytes; b++) n_exc += (u32)__builtin_popcount(rd8(lx, L.exceptionMask + b));
    const u32 exc_off = rd32(lx, limex::exceptionOffset);
    if (n_exc != rd32(lx, limex::exceptionCount) || !lx_range(exc_off, (size_t)n_exc * L.E_SIZE)) { why = "LimEx exception table inconsistent"; return E_INVALID; }
    for (u32 i = 0; i < n_exc; i++) {
        const u8 *e = lx + exc_off + (size_t)i * L.E_SIZE;
        if (rd8(e, L.e_trigger)) { why = "LimEx repeat triggers are not supported yet"; return E_ARCH; }
        const u8 sq = rd8(e, L.e_hasSquash);
        if (sq != limex::SQUASH_NONE && sq != limex::SQUASH_CYCLIC && sq != limex::SQUASH_REPORT) { why = "LimEx tug squash without repeats"; return E_INVALID; }
        const u32 rep = rd32(e, L.e_reports);
        if (rep != 0xffffffffu && !add_list(rep)) { why = "LimEx report list outside the engine"; return E_INVALID; }
    }
    // accept tables: one NFAAccept per bit of accept / acceptAtEOD
    const size_t masks[2] = {L.accept, L.acceptAtEOD}, cnt[2] = {limex::acceptCount, limex::acceptEodCount},
                 offs[2] = {limex::acceptOffset, limex::acceptEodOffset};
    for (int t = 0; t < 2; t++) {
        u32 na = 0;
        for (u32 b = 0; b < L.bytes; b++) na += (u32)__builtin_popcount(rd8(lx, masks[t] + b));
        const u32 ao = rd32(lx, offs[t]);
        if (na != rd32(lx, cnt[t]) || !lx_range(ao, (size_t)na * nfaaccept::SIZE)) { why = "LimEx accept table inconsistent"; return E_INVALID; }
        for (u32 i = 0; i < na; i++) {
            const u8 *a = lx + ao + (size_t)i * nfaaccept::SIZE;
            if (rd8(a, nfaaccept::single_report)) progs.insert(rd32(a, nfaaccept::reports));
            else if (!add_list(rd32(a, nfaaccept::reports))) { why = "LimEx accept report list outside the engine"; return E_INVALID; }
        }
    }
    d.max_lit_len = 8;
    for (u32 p : progs) {
        if (p < rose::SIZE || !in_range(d, p, 8)) { why = "report program offset outside bytecode"; return E_INVALID; }
        int rv = census_program(d, p, why, d.max_lit_len);
        if (rv != E_SUCCESS) return rv;
        d.programs.push_back(p);
    }
    return E_SUCCESS;
}

// `db` is the caller's in-memory hs_database_t.
inline int parse_database(const void *db, DbView &d, std::string &why) {
    if (!db) return E_INVALID;
    if (rd32(db, db::magic) != db::MAGIC) { why = "bad magic"; return E_INVALID; }
    if (rd32(db, db::version) != db::VERSION_5_4_11) { why = "database version mismatch"; return E_DB_VERSION; }
    const u32 bc_off = rd32(db, db::bytecode);
    d.bc = reinterpret_cast<const u8 *>(db) + bc_off;
    d.bc_len = rd32(db, db::length);
    d.crc = rd32(db, db::crc32);
    if ((reinterpret_cast<uintptr_t>(d.bc) & 15) != 0) { why = "bytecode not 16-byte aligned"; return E_INVALID; }
    if (d.bc_len < rose::SIZE) { why = "bytecode too short"; return E_INVALID; }
    const u8 *r = d.bc;
    const u32 mode = rd32(r, rose::mode);
    if (mode != MODE_BLOCK && mode != MODE_VECTORED && mode != MODE_STREAM) { why = "unknown database mode"; return E_DB_MODE; }
    d.vectored = mode == MODE_VECTORED;
    d.streaming = mode == MODE_STREAM;
    d.history_required = rd32(r, rose::historyRequired);
    // stream state that the device path carries: pure-literal databases (history, exhaustion: on the host) and, in stream
    // mode, ONE DFA engine (its state id per stream goes into and comes out of every block scan)
    if ((d.vectored && rd8(r, rose::runtimeImpl) != rose::RUNTIME_PURE_LITERAL) ||
        (d.streaming && rd8(r, rose::runtimeImpl) == rose::RUNTIME_FULL_ROSE)) {
        why = "vectored mode is served for pure-literal databases, stream mode for pure-literal and single-DFA databases only";
        return E_ARCH;
    }
    if (rd32(r, rose::size) > d.bc_len) { why = "engine size exceeds database length"; return E_INVALID; }
    if (rd8(r, rose::hasSom)) { why = "start-of-match databases are not supported"; return E_ARCH; }
    if (rd32(r, rose::ckeyCount) || rd32(r, rose::lkeyCount)) { why = "logical combinations are not supported"; return E_ARCH; }
    const u8 impl = rd8(r, rose::runtimeImpl);
    if (impl != rose::RUNTIME_FULL_ROSE && impl != rose::RUNTIME_PURE_LITERAL && impl != rose::RUNTIME_SINGLE_OUTFIX) { why = "unknown runtimeImpl"; return E_INVALID; }
    const u8 *bd = r + rose::boundary;
    if (rd32(bd, rose::bd::reportEodOffset) || rd32(bd, rose::bd::reportZeroOffset) ||
        rd32(bd, rose::bd::reportZeroEodOffset)) {
        why = "boundary (vacuous-pattern) report programs are not supported yet";
        return E_ARCH;
    }
    d.min_width = rd32(r, rose::minWidth);
    d.min_width_excl = rd32(r, rose::minWidthExcludingBoundaries);
    d.max_bi_width = rd32(r, rose::maxBiAnchoredWidth);
    // the stream path has no length gates (hs_scan_stream_internal, src/runtime.c:880-975, tests none of them)
    if (d.vectored || d.streaming) { d.min_width = d.min_width_excl = 0; d.max_bi_width = rose::BOUND_INF; }
    d.ekey_count = rd32(r, rose::ekeyCount);
    d.dkey_count = rd32(r, rose::dkeyCount);
    d.groups = rd64(r, rose::initialGroups) & rd64(r, rose::floating_group_mask);
    if (impl == rose::RUNTIME_SINGLE_OUTFIX) {
        const int orv = parse_single_outfix(d, why);
        if (orv != E_SUCCESS || !d.streaming) return orv;
        if (d.kind != ENG_MCCLELLAN && d.kind != ENG_SHENG) { why = "stream mode: only DFA engines carry their state across blocks on the device path"; return E_ARCH; }
        // (end-of-data accepts are raised by hs_gpu_close_stream: needs requiresEodCheck, as the reference's close does)
        if ((d.nfa_flags & nfa::ACCEPTS_EOD) && !rd8(r, rose::requiresEodCheck)) { why = "stream mode: engine accepts at end of data but the header does not ask for the check"; return E_INVALID; }
        return E_SUCCESS;
    }
    if (impl == rose::RUNTIME_FULL_ROSE) {
        // supported shapes: outfix engines only; a floating literal matcher (programs through the full interpreter,
        // roseFloatingCallback src/rose/match.c:479-523); both (literal matches and engine matches merge by offset)
        const int grv = full_rose_gates(d, why);
        if (grv != E_SUCCESS) return grv;
        if (!rd32(r, rose::fmatcherOffset)) return parse_outfix_rose(d, why);
    }
    d.hwlm_off = rd32(r, rose::fmatcherOffset);
    if (!d.hwlm_off || !in_range(d, d.hwlm_off, hwlm::BODY + 24)) { why = "no floating literal matcher"; return E_INVALID; }
    const u8 type = rd8(r, d.hwlm_off + hwlm::type);
    d.eng_off = d.hwlm_off + (u32)hwlm::BODY;
    std::set<u32> progs;
    if (type == hwlm::ENGINE_NOOD) {
        if (!in_range(d, d.eng_off, nood::SIZE)) { why = "noodle table truncated"; return E_INVALID; }
        d.kind = ENG_NOODLE;
        d.nood_id = rd32(r, d.eng_off + nood::id);
        d.nood_msk = rd64(r, d.eng_off + nood::msk);
        d.nood_cmp = rd64(r, d.eng_off + nood::cmp);
        d.nood_len = rd8(r, d.eng_off + nood::msk_len);
        if (d.nood_len < 1 || d.nood_len > 8) { why = "noodle mask length out of range"; return E_INVALID; }
        d.max_string_len = d.nood_len;
        d.n_strings = 1;
        progs.insert(d.nood_id);
    } else if (type == hwlm::ENGINE_FDR) {
        const u8 *e = r + d.eng_off;
        d.engine_id = rd32(e, fdr::engineID);
        const u32 esize = rd32(e, fdr::size);
        if (!in_range(d, d.eng_off, esize) || esize < fdr::TABLE) { why = "literal matcher truncated"; return E_INVALID; }
        d.max_string_len = rd32(e, fdr::maxStringLen);
        d.n_strings = rd32(e, fdr::numStrings);
        d.conf_base = d.eng_off + rd32(e, fdr::confOffset);
        if (d.engine_id == 0) {
            d.kind = ENG_FDR;
            d.fdr_stride = rd8(e, fdr::stride);
            d.fdr_domain = rd8(e, fdr::domain);
            d.fdr_tab = d.eng_off + (u32)fdr::TABLE;
            d.fdr_start = rd64(e, fdr::start);
            if (d.fdr_domain < 9 || d.fdr_domain > 15 || (d.fdr_stride != 1 && d.fdr_stride != 2 && d.fdr_stride != 4) ||
                rd32(e, fdr::tabSize) != (8u << d.fdr_domain) || !in_range(d, d.fdr_tab, (size_t)8 << d.fdr_domain)) {
                why = "malformed FDR header";
                return E_INVALID;
            }
        } else if (d.engine_id >= teddy::ENGINE_FIRST && d.engine_id <= teddy::ENGINE_LAST) {
            d.kind = ENG_TEDDY;
            d.teddy_masks = (d.engine_id - teddy::ENGINE_FIRST) / 2 + 1;
            d.teddy_mask_off = d.eng_off + (u32)teddy::MASKS;
            if (!in_range(d, d.teddy_mask_off, (size_t)d.teddy_masks * 32)) { why = "Teddy masks truncated"; return E_INVALID; }
        } else if (d.engine_id >= teddy::FAT_FIRST && d.engine_id <= teddy::FAT_LAST) {
            // fat Teddy (AVX2 hosts, src/fdr/teddy_avx2.c:452-480): 16 buckets, every nibble mask twice -- buckets 0-7 then
            // buckets 8-15 (teddy_compile.cpp:436-500, maskWidth = 2); confBase has 16 entries
            d.kind = ENG_TEDDY;
            d.teddy_fat = 1;
            d.n_buckets = 16;
            d.teddy_masks = (d.engine_id - teddy::FAT_FIRST) / 2 + 1;
            d.teddy_mask_off = d.eng_off + (u32)teddy::MASKS;
            if (!in_range(d, d.teddy_mask_off, (size_t)d.teddy_masks * 64)) { why = "fat Teddy masks truncated"; return E_INVALID; }
        } else {
            why = "unknown literal matcher engine id";
            return E_INVALID;
        }
        if (!in_range(d, d.conf_base, d.n_buckets * 4)) { why = "confirm base outside bytecode"; return E_INVALID; }
        for (u32 b = 0; b < d.n_buckets; b++) {
            const u32 cf = rd32(r, d.conf_base + 4 * b);
            if (!cf) continue;
            const size_t fc = (size_t)d.conf_base + cf;
            if (!in_range(d, fc, conf::SIZE)) { why = "FDRConfirm outside bytecode"; return E_INVALID; }
            const u32 nbits = rd32(r, fc + conf::nBits);
            if (nbits > 24 || !in_range(d, fc + conf::SIZE, (size_t)4 << nbits)) { why = "confirm index outside bytecode"; return E_INVALID; }
            for (u32 h = 0; h < (1u << nbits); h++) {
                u32 start = rd32(r, fc + conf::SIZE + 4 * (size_t)h);
                if (!start) continue;
                size_t li = fc + start;
                for (;;) {
                    if (!in_range(d, li, lit::SIZE)) { why = "LitInfo outside bytecode"; return E_INVALID; }
                    LitRec lr;
                    lr.v = rd64(r, li + lit::v);
                    lr.msk = rd64(r, li + lit::msk);
                    lr.groups = rd64(r, li + lit::groups);
                    lr.program = rd32(r, li + lit::id);
                    lr.size = rd8(r, li + lit::size);
                    lr.flags = rd8(r, li + lit::flags);
                    lr.bucket = (u8)b;
                    d.lits.push_back(lr);
                    progs.insert(lr.program);
                    if (!rd8(r, li + lit::next)) break;
                    li += lit::SIZE;
                }
            }
        }
    } else {
        why = "unknown HWLM engine type";
        return E_INVALID;
    }
    d.max_lit_len = 8;
    for (u32 p : progs) {
        if (p < rose::SIZE || !in_range(d, p, 8)) { why = "literal program offset outside bytecode"; return E_INVALID; }
        int rv = census_program(d, p, why, d.max_lit_len);
        if (rv != E_SUCCESS) return rv;
        d.programs.push_back(p);
    }
    if (d.streaming) {
        // one exhaustion key per external id, and no silent exhaustion: the host carries a stream's exhausted keys from
        // call to call and sees only what is reported
        u64 hist[128] = {0};
        std::vector<std::pair<u32, u32>> reps;
        u32 dummy = 0;
        for (u32 p : d.programs) census_program(d, p, why, dummy, hist, &reps);
        if (hist[op::SET_EXHAUST]) { why = "stream mode: quiet exhaustion (SET_EXHAUST) is not supported"; return E_ARCH; }
        std::sort(reps.begin(), reps.end());
        reps.erase(std::unique(reps.begin(), reps.end()), reps.end());
        for (size_t k = 1; k < reps.size(); k++)
            if (reps[k].first == reps[k - 1].first) { why = "stream mode: report id " + std::to_string(reps[k].first) + " has more than one exhaustion key"; return E_ARCH; }
        d.stream_ekey_of = reps;
        for (auto &e : reps) if (e.second != ~0u) d.stream_any_ekey = true;
    }
    if (impl == rose::RUNTIME_FULL_ROSE && rd32(r, rose::queueCount)) return parse_outfix_queues(d, why);
    return E_SUCCESS;
}

}  // namespace hsb
