// tests/layout_dump.cpp -- prints the constants of vectorscan_b200/csrc/hsb_layout.h under the key names that
// oracle/layout_probe.c prints for the reference's own structs; tests/test_layout.py diffs the two.
#include "hsb_layout.h"
#include <cstdio>
using namespace hsb;
#define P(key, val) std::printf("%s %lld\n", key, (long long)(val))
#define R(f) P("struct RoseEngine." #f, rose::f)
#define SO(f) P("struct RoseStateOffsets." #f, rose::so::f)
#define OPF(ns, f) P("op." #ns "." #f, op::ns::f)
int main() {
    P("sizeof.struct hs_database", db::SIZE); P("struct hs_database.magic", db::magic);
    P("struct hs_database.version", db::version); P("struct hs_database.length", db::length);
    P("struct hs_database.platform", db::platform); P("struct hs_database.crc32", db::crc32);
    P("struct hs_database.reserved0", db::reserved0); P("struct hs_database.reserved1", db::reserved1);
    P("struct hs_database.bytecode", db::bytecode); P("struct hs_database.padding", db::padding);
    P("struct hs_database.bytes", db::bytes);
    P("HS_DB_VERSION", db::VERSION_5_4_11); P("HS_DB_MAGIC", db::MAGIC);
    P("HS_PLATFORM_NOAVX2", db::NOAVX2); P("HS_PLATFORM_NOAVX512", db::NOAVX512); P("HS_PLATFORM_NOAVX512VBMI", db::NOAVX512VBMI);
    P("HS_MODE_BLOCK", MODE_BLOCK); P("HS_MODE_STREAM", MODE_STREAM); P("HS_MODE_VECTORED", MODE_VECTORED);
    P("sizeof.struct RoseEngine", rose::SIZE);
    R(pureLiteral); R(noFloatingRoots); R(requiresEodCheck); R(hasOutfixesInSmallBlock); R(runtimeImpl);
    R(mpvTriggeredByLeaf); R(canExhaust); R(hasSom); R(somHorizon); R(mode); R(historyRequired); R(ekeyCount);
    R(lkeyCount); R(lopCount); R(ckeyCount); R(logicalTreeOffset); R(combInfoMapOffset); R(dkeyCount); R(dkeyLogSize);
    R(invDkeyOffset); R(somLocationCount); R(somLocationFatbitSize); R(rolesWithStateCount); R(stateSize);
    R(anchorStateSize); R(tStateSize); R(scratchStateSize); R(smallWriteOffset); R(amatcherOffset); R(ematcherOffset);
    R(fmatcherOffset); R(drmatcherOffset); R(sbmatcherOffset); R(longLitTableOffset); R(amatcherMinWidth);
    R(fmatcherMinWidth); R(eodmatcherMinWidth); R(amatcherMaxBiAnchoredWidth); R(fmatcherMaxBiAnchoredWidth);
    R(reportProgramOffset); R(reportProgramCount); R(delayProgramOffset); R(anchoredProgramOffset);
    R(activeArrayCount); R(activeLeftCount); R(queueCount); R(activeQueueArraySize); R(eagerIterOffset);
    R(handledKeyCount); R(handledKeyFatbitSize); R(leftOffset); R(roseCount); R(eodProgramOffset);
    R(flushCombProgramOffset); R(lastFlushCombProgramOffset); R(lastByteHistoryIterOffset); R(minWidth);
    R(minWidthExcludingBoundaries); R(maxBiAnchoredWidth); R(anchoredDistance); R(anchoredMinDistance);
    R(floatingDistance); R(floatingMinDistance); R(smallBlockDistance); R(floatingMinLiteralMatchOffset);
    R(nfaInfoOffset); R(initialGroups); R(floating_group_mask); R(size); R(delay_count); R(delay_fatbit_size);
    R(anchored_count); R(anchored_fatbit_size); R(maxFloatingDelayedMatch); R(delayRebuildLength); R(stateOffsets);
    R(boundary); R(totalNumLiterals); R(asize); R(outfixBeginQueue); R(outfixEndQueue); R(leftfixBeginQueue);
    R(initMpvNfa); R(rosePrefixCount); R(activeLeftIterOffset); R(ematcherRegionSize); R(somRevCount);
    R(somRevOffsetOffset); R(longLitStreamState); R(state_init);
    P("sizeof.struct RoseStateOffsets", rose::so::SIZE);
    SO(history); SO(exhausted); SO(exhausted_size); SO(logicalVec); SO(logicalVec_size); SO(combVec); SO(combVec_size);
    SO(activeLeafArray); SO(activeLeafArray_size); SO(activeLeftArray); SO(activeLeftArray_size); SO(leftfixLagTable);
    SO(anchorState); SO(groups); SO(groups_size); SO(longLitState); SO(longLitState_size); SO(somLocation);
    SO(somValid); SO(somWritable); SO(somMultibit_size); SO(nfaStateBegin); SO(end);
    P("struct RoseBoundaryReports.reportEodOffset", rose::bd::reportEodOffset);
    P("struct RoseBoundaryReports.reportZeroOffset", rose::bd::reportZeroOffset);
    P("struct RoseBoundaryReports.reportZeroEodOffset", rose::bd::reportZeroEodOffset);
    P("ROSE_RUNTIME_FULL_ROSE", rose::RUNTIME_FULL_ROSE); P("ROSE_RUNTIME_PURE_LITERAL", rose::RUNTIME_PURE_LITERAL);
    P("ROSE_RUNTIME_SINGLE_OUTFIX", rose::RUNTIME_SINGLE_OUTFIX); P("ROSE_BOUND_INF", rose::BOUND_INF);
    P("sizeof.struct HWLM", hwlm::SIZE); P("struct HWLM.type", hwlm::type); P("struct HWLM.accel1_groups", hwlm::accel1_groups);
    P("struct HWLM.accel1", hwlm::accel1); P("struct HWLM.accel0", hwlm::accel0);
    P("HWLM_ENGINE_NOOD", hwlm::ENGINE_NOOD); P("HWLM_ENGINE_FDR", hwlm::ENGINE_FDR); P("sizeof.union AccelAux", hwlm::ACCEL_AUX_SIZE);
    P("ACCEL_NONE", accel::NONE); P("ACCEL_VERM", accel::VERM); P("ACCEL_VERM_NOCASE", accel::VERM_NOCASE);
    P("ACCEL_DVERM", accel::DVERM); P("ACCEL_DVERM_NOCASE", accel::DVERM_NOCASE); P("ACCEL_SHUFTI", accel::SHUFTI);
    P("ACCEL_DSHUFTI", accel::DSHUFTI); P("ACCEL_TRUFFLE", accel::TRUFFLE); P("ACCEL_RED_TAPE", accel::RED_TAPE);
    P("ACCEL_DVERM_MASKED", accel::DVERM_MASKED);
    P("accel.offset", accel::offset); P("accel.verm_c", accel::verm_c); P("accel.dverm_c1", accel::dverm_c1);
    P("accel.dverm_c2", accel::dverm_c2); P("accel.dverm_m1", accel::dverm_m1); P("accel.dverm_m2", accel::dverm_m2);
    P("accel.shufti_lo", accel::shufti_lo); P("accel.shufti_hi", accel::shufti_hi);
    P("accel.truffle_mask1", accel::truffle_mask1); P("accel.truffle_mask2", accel::truffle_mask2);
    P("sizeof.struct noodTable", nood::SIZE); P("struct noodTable.id", nood::id); P("struct noodTable.msk", nood::msk);
    P("struct noodTable.cmp", nood::cmp); P("struct noodTable.msk_len", nood::msk_len);
    P("struct noodTable.key_offset", nood::key_offset); P("struct noodTable.nocase", nood::nocase);
    P("struct noodTable.single", nood::single); P("struct noodTable.key0", nood::key0); P("struct noodTable.key1", nood::key1);
    P("sizeof.struct FDR", fdr::SIZE); P("struct FDR.engineID", fdr::engineID); P("struct FDR.size", fdr::size);
    P("struct FDR.maxStringLen", fdr::maxStringLen); P("struct FDR.numStrings", fdr::numStrings);
    P("struct FDR.confOffset", fdr::confOffset); P("struct FDR.floodOffset", fdr::floodOffset);
    P("struct FDR.stride", fdr::stride); P("struct FDR.domain", fdr::domain); P("struct FDR.domainMask", fdr::domainMask);
    P("struct FDR.tabSize", fdr::tabSize); P("struct FDR.start", fdr::start);
    P("sizeof.struct Teddy", teddy::SIZE);
    P("sizeof.struct FDRFlood", flood::SIZE); P("struct FDRFlood.allGroups", flood::allGroups);
    P("struct FDRFlood.suffix", flood::suffix); P("struct FDRFlood.idCount", flood::idCount);
    P("struct FDRFlood.ids", flood::ids); P("struct FDRFlood.groups", flood::groups); P("FDR_FLOOD_MAX_IDS", flood::MAX_IDS);
    P("sizeof.struct FDRConfirm", conf::SIZE); P("struct FDRConfirm.andmsk", conf::andmsk); P("struct FDRConfirm.mult", conf::mult);
    P("struct FDRConfirm.nBits", conf::nBits); P("struct FDRConfirm.groups", conf::groups);
    P("sizeof.struct LitInfo", lit::SIZE); P("struct LitInfo.v", lit::v); P("struct LitInfo.msk", lit::msk);
    P("struct LitInfo.groups", lit::groups); P("struct LitInfo.id", lit::id); P("struct LitInfo.size", lit::size);
    P("struct LitInfo.flags", lit::flags); P("struct LitInfo.next", lit::next); P("FDR_LIT_FLAG_NOREPEAT", lit::FLAG_NOREPEAT);
    OPF(check_groups, groups); OPF(check_groups, SIZE);
    OPF(check_mask, and_mask); OPF(check_mask, cmp_mask); OPF(check_mask, neg_mask); OPF(check_mask, offset);
    OPF(check_mask, fail_jump); OPF(check_mask, SIZE);
    OPF(check_mask32, and_mask); OPF(check_mask32, cmp_mask); OPF(check_mask32, neg_mask); OPF(check_mask32, offset);
    OPF(check_mask32, fail_jump); OPF(check_mask32, SIZE);
    OPF(check_byte, and_mask); OPF(check_byte, cmp_mask); OPF(check_byte, negation); OPF(check_byte, offset);
    OPF(check_byte, fail_jump); OPF(check_byte, SIZE);
    OPF(check_shufti_16x8, nib_mask); OPF(check_shufti_16x8, bucket_select_mask); OPF(check_shufti_16x8, neg_mask);
    OPF(check_shufti_16x8, offset); OPF(check_shufti_16x8, fail_jump); OPF(check_shufti_16x8, SIZE);
    OPF(check_shufti_32x8, hi_mask); OPF(check_shufti_32x8, lo_mask); OPF(check_shufti_32x8, bucket_select_mask);
    OPF(check_shufti_32x8, neg_mask); OPF(check_shufti_32x8, offset); OPF(check_shufti_32x8, fail_jump); OPF(check_shufti_32x8, SIZE);
    OPF(check_shufti_16x16, hi_mask); OPF(check_shufti_16x16, lo_mask); OPF(check_shufti_16x16, bucket_select_mask);
    OPF(check_shufti_16x16, neg_mask); OPF(check_shufti_16x16, offset); OPF(check_shufti_16x16, fail_jump); OPF(check_shufti_16x16, SIZE);
    OPF(check_shufti_32x16, hi_mask); OPF(check_shufti_32x16, lo_mask); OPF(check_shufti_32x16, bucket_select_mask_hi);
    OPF(check_shufti_32x16, bucket_select_mask_lo); OPF(check_shufti_32x16, neg_mask); OPF(check_shufti_32x16, offset);
    OPF(check_shufti_32x16, fail_jump); OPF(check_shufti_32x16, SIZE);
    OPF(dedupe, quash_som); OPF(dedupe, dkey); OPF(dedupe, offset_adjust); OPF(dedupe, fail_jump); OPF(dedupe, SIZE);
    OPF(report, onmatch); OPF(report, offset_adjust); OPF(report, SIZE);
    P("op.final_report.onmatch", op::report::onmatch); P("op.final_report.offset_adjust", op::report::offset_adjust);
    P("op.final_report.SIZE", op::report::SIZE);
    OPF(report_exhaust, onmatch); OPF(report_exhaust, offset_adjust); OPF(report_exhaust, ekey); OPF(report_exhaust, SIZE);
    OPF(dedupe_and_report, quash_som); OPF(dedupe_and_report, dkey); OPF(dedupe_and_report, onmatch);
    OPF(dedupe_and_report, offset_adjust); OPF(dedupe_and_report, fail_jump); OPF(dedupe_and_report, SIZE);
    OPF(check_exhausted, ekey); OPF(check_exhausted, fail_jump); OPF(check_exhausted, SIZE);
    OPF(squash_groups, groups); OPF(squash_groups, SIZE);
    OPF(check_lit, lit_offset); OPF(check_lit, lit_length); OPF(check_lit, fail_jump); OPF(check_lit, SIZE);
    OPF(included_jump, squash); OPF(included_jump, child_offset); OPF(included_jump, SIZE);
    OPF(set_exhaust, ekey); OPF(set_exhaust, SIZE);
    P("op.end.SIZE", op::SIZE_1);
    P("ROSE_INSTR_END", op::END); P("ROSE_INSTR_CHECK_GROUPS", op::CHECK_GROUPS); P("ROSE_INSTR_CHECK_MASK", op::CHECK_MASK);
    P("ROSE_INSTR_CHECK_MASK_32", op::CHECK_MASK_32); P("ROSE_INSTR_CHECK_BYTE", op::CHECK_BYTE);
    P("ROSE_INSTR_CHECK_SHUFTI_16x8", op::CHECK_SHUFTI_16x8); P("ROSE_INSTR_DEDUPE", op::DEDUPE);
    P("ROSE_INSTR_REPORT", op::REPORT); P("ROSE_INSTR_REPORT_EXHAUST", op::REPORT_EXHAUST);
    P("ROSE_INSTR_DEDUPE_AND_REPORT", op::DEDUPE_AND_REPORT); P("ROSE_INSTR_FINAL_REPORT", op::FINAL_REPORT);
    P("ROSE_INSTR_CHECK_EXHAUSTED", op::CHECK_EXHAUSTED); P("ROSE_INSTR_SQUASH_GROUPS", op::SQUASH_GROUPS);
    P("ROSE_INSTR_CHECK_LONG_LIT", op::CHECK_LONG_LIT); P("ROSE_INSTR_CHECK_MED_LIT", op::CHECK_MED_LIT);
    P("ROSE_INSTR_CHECK_MED_LIT_NOCASE", op::CHECK_MED_LIT_NOCASE); P("ROSE_INSTR_CLEAR_WORK_DONE", op::CLEAR_WORK_DONE);
    P("ROSE_INSTR_INCLUDED_JUMP", op::INCLUDED_JUMP); P("ROSE_INSTR_SET_EXHAUST", op::SET_EXHAUST);
    P("ROSE_INSTR_CHECK_SHUFTI_64x8", op::CHECK_SHUFTI_64x8); P("ROSE_INSTR_CHECK_MASK_64", op::CHECK_MASK_64);
    P("LAST_ROSE_INSTRUCTION", op::LAST);
    // NFA engines
    P("sizeof.struct NFA", nfa::SIZE); P("struct NFA.flags", nfa::flags); P("struct NFA.length", nfa::length);
    P("struct NFA.type", nfa::type); P("struct NFA.rAccelType", nfa::rAccelType); P("struct NFA.rAccelOffset", nfa::rAccelOffset);
    P("struct NFA.maxBiAnchoredWidth", nfa::maxBiAnchoredWidth); P("struct NFA.rAccelData", nfa::rAccelData);
    P("struct NFA.queueIndex", nfa::queueIndex); P("struct NFA.nPositions", nfa::nPositions);
    P("struct NFA.scratchStateSize", nfa::scratchStateSize); P("struct NFA.streamStateSize", nfa::streamStateSize);
    P("struct NFA.maxWidth", nfa::maxWidth); P("struct NFA.minWidth", nfa::minWidth); P("struct NFA.maxOffset", nfa::maxOffset);
    P("LIMEX_NFA_32", nfa::LIMEX_32); P("LIMEX_NFA_64", nfa::LIMEX_64); P("LIMEX_NFA_128", nfa::LIMEX_128);
    P("LIMEX_NFA_256", nfa::LIMEX_256); P("LIMEX_NFA_384", nfa::LIMEX_384); P("LIMEX_NFA_512", nfa::LIMEX_512);
    P("NFA_ACCEPTS_EOD", nfa::ACCEPTS_EOD); P("NFA_ZOMBIE", nfa::ZOMBIE); P("MO_INVALID_IDX", INVALID_DKEY);
    for (u32 t = 0; t < 6; t++) {
        const limex::Layout l = limex::layout(t);
        char k[96];
#define LK(f, v) std::snprintf(k, sizeof k, "struct LimExNFA%u." #f, l.bits); P(k, v)
#define EK(f, v) std::snprintf(k, sizeof k, "struct NFAException%u." #f, l.bits); P(k, v)
        std::snprintf(k, sizeof k, "sizeof.struct LimExNFA%u", l.bits); P(k, l.SIZE);
        LK(reachMap, limex::reachMap); LK(reachSize, limex::reachSize); LK(accelCount, limex::accelCount);
        LK(accelTableOffset, limex::accelTableOffset); LK(accelAuxCount, limex::accelAuxCount);
        LK(accelAuxOffset, limex::accelAuxOffset); LK(acceptCount, limex::acceptCount); LK(acceptOffset, limex::acceptOffset);
        LK(acceptEodCount, limex::acceptEodCount); LK(acceptEodOffset, limex::acceptEodOffset);
        LK(exceptionCount, limex::exceptionCount); LK(exceptionOffset, limex::exceptionOffset);
        LK(repeatCount, limex::repeatCount); LK(repeatOffset, limex::repeatOffset); LK(squashOffset, limex::squashOffset);
        LK(squashCount, limex::squashCount); LK(topCount, limex::topCount); LK(topOffset, limex::topOffset);
        LK(stateSize, limex::stateSize); LK(flags, limex::flags);
        LK(init, l.init); LK(initDS, l.initDS); LK(accept, l.accept); LK(acceptAtEOD, l.acceptAtEOD); LK(accel, l.accel);
        LK(accelPermute, l.accelPermute); LK(accelCompare, l.accelCompare); LK(accel_and_friends, l.accel_and_friends);
        LK(compressMask, l.compressMask); LK(exceptionMask, l.exceptionMask); LK(repeatCyclicMask, l.repeatCyclicMask);
        LK(zombieMask, l.zombieMask); LK(shift, l.shift); LK(shiftCount, l.shiftCount); LK(shiftAmount, l.shiftAmount);
        LK(exceptionShufMask, l.exceptionShufMask); LK(exceptionBitMask, l.exceptionBitMask);
        LK(exceptionAndMask, l.exceptionAndMask);
        std::snprintf(k, sizeof k, "sizeof.struct NFAException%u", l.bits); P(k, l.E_SIZE);
        EK(squash, l.e_squash); EK(successors, l.e_successors); EK(reports, l.e_reports); EK(repeatOffset, l.e_repeatOffset);
        EK(hasSquash, l.e_hasSquash); EK(trigger, l.e_trigger);
    }
    P("sizeof.struct NFAAccept", nfaaccept::SIZE); P("struct NFAAccept.single_report", nfaaccept::single_report);
    P("struct NFAAccept.reports", nfaaccept::reports); P("struct NFAAccept.squash", nfaaccept::squash);
    P("sizeof.struct NfaInfo", nfainfo::SIZE); P("struct NfaInfo.nfaOffset", nfainfo::nfaOffset);
    P("struct NfaInfo.stateOffset", nfainfo::stateOffset); P("struct NfaInfo.fullStateOffset", nfainfo::fullStateOffset);
    P("struct NfaInfo.no_retrigger", nfainfo::no_retrigger); P("struct NfaInfo.in_sbmatcher", nfainfo::in_sbmatcher);
    P("struct NfaInfo.eod", nfainfo::eod);
    return 0;
}
