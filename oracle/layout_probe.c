/* oracle/layout_probe.c -- TEST INFRASTRUCTURE ONLY.
 * Prints sizeof/offsetof facts about the reference's bytecode structs, taken from the reference's
 * own headers, as "name value" lines.  tests/test_layout.py compares them with the constants the
 * product declares in vectorscan_b200/csrc/hsb_layout.h.  */
#include "config.h"
#include <stdio.h>
#include <stddef.h>
#include "ue2common.h"
#include "database.h"
#include "scratch.h"
#include "rose/rose_internal.h"
#include "rose/rose_program.h"
#include "hwlm/hwlm_internal.h"
#include "hwlm/noodle_internal.h"
#include "fdr/fdr_internal.h"
#include "fdr/fdr_confirm.h"
#include "fdr/teddy_internal.h"
#include "nfa/nfa_internal.h"
#include "nfa/limex_internal.h"
#include "nfa/accel.h"
#include "nfa/mcclellan_internal.h"
#define report_list report_list_sheng /* both engine headers define struct report_list */
#include "nfa/sheng_internal.h"
#undef report_list
#define report_list report_list_mcsheng /* mcsheng_internal.h repeats McClellan's helper structs */
#define mstate_aux mstate_aux_mcsheng
#include "nfa/mcsheng_internal.h"
#undef report_list
#undef mstate_aux
#include "util/multibit_internal.h"
#include "hs_compile.h"

#define SZ(t) printf("sizeof." #t " %zu\n", sizeof(t))
#define OF(t, f) printf(#t "." #f " %zu\n", offsetof(t, f))
#define VAL(v) printf(#v " %lld\n", (long long)(v))
#define LIMEX(n) \
 SZ(struct LimExNFA##n); OF(struct LimExNFA##n, reachMap); OF(struct LimExNFA##n, reachSize); OF(struct LimExNFA##n, accelCount); \
 OF(struct LimExNFA##n, accelTableOffset); OF(struct LimExNFA##n, accelAuxCount); OF(struct LimExNFA##n, accelAuxOffset); \
 OF(struct LimExNFA##n, acceptCount); OF(struct LimExNFA##n, acceptOffset); OF(struct LimExNFA##n, acceptEodCount); OF(struct LimExNFA##n, acceptEodOffset); \
 OF(struct LimExNFA##n, exceptionCount); OF(struct LimExNFA##n, exceptionOffset); OF(struct LimExNFA##n, repeatCount); OF(struct LimExNFA##n, repeatOffset); \
 OF(struct LimExNFA##n, squashOffset); OF(struct LimExNFA##n, squashCount); OF(struct LimExNFA##n, topCount); OF(struct LimExNFA##n, topOffset); \
 OF(struct LimExNFA##n, stateSize); OF(struct LimExNFA##n, flags); OF(struct LimExNFA##n, init); OF(struct LimExNFA##n, initDS); OF(struct LimExNFA##n, accept); \
 OF(struct LimExNFA##n, acceptAtEOD); OF(struct LimExNFA##n, accel); OF(struct LimExNFA##n, accelPermute); OF(struct LimExNFA##n, accelCompare); \
 OF(struct LimExNFA##n, accel_and_friends); OF(struct LimExNFA##n, compressMask); OF(struct LimExNFA##n, exceptionMask); OF(struct LimExNFA##n, repeatCyclicMask); \
 OF(struct LimExNFA##n, zombieMask); OF(struct LimExNFA##n, shift); OF(struct LimExNFA##n, shiftCount); OF(struct LimExNFA##n, shiftAmount); \
 OF(struct LimExNFA##n, exceptionShufMask); OF(struct LimExNFA##n, exceptionBitMask); OF(struct LimExNFA##n, exceptionAndMask); \
 SZ(struct NFAException##n); OF(struct NFAException##n, squash); OF(struct NFAException##n, successors); OF(struct NFAException##n, reports); \
 OF(struct NFAException##n, repeatOffset); OF(struct NFAException##n, hasSquash); OF(struct NFAException##n, trigger);

int main(void) {
    SZ(struct hs_database);
    OF(struct hs_database, magic); OF(struct hs_database, version); OF(struct hs_database, length);
    OF(struct hs_database, platform); OF(struct hs_database, crc32); OF(struct hs_database, bytecode);
    OF(struct hs_database, bytes);
    VAL(HS_DB_VERSION); VAL(HS_DB_MAGIC);
    VAL(HS_PLATFORM_NOAVX2); VAL(HS_PLATFORM_NOAVX512); VAL(HS_PLATFORM_NOAVX512VBMI);

    SZ(struct RoseEngine);
    OF(struct RoseEngine, pureLiteral); OF(struct RoseEngine, noFloatingRoots);
    OF(struct RoseEngine, requiresEodCheck); OF(struct RoseEngine, hasOutfixesInSmallBlock);
    OF(struct RoseEngine, runtimeImpl); OF(struct RoseEngine, mpvTriggeredByLeaf);
    OF(struct RoseEngine, canExhaust); OF(struct RoseEngine, hasSom); OF(struct RoseEngine, somHorizon);
    OF(struct RoseEngine, mode); OF(struct RoseEngine, historyRequired); OF(struct RoseEngine, ekeyCount);
    OF(struct RoseEngine, lkeyCount); OF(struct RoseEngine, lopCount); OF(struct RoseEngine, ckeyCount);
    OF(struct RoseEngine, dkeyCount); OF(struct RoseEngine, dkeyLogSize); OF(struct RoseEngine, invDkeyOffset);
    OF(struct RoseEngine, somLocationCount); OF(struct RoseEngine, somLocationFatbitSize);
    OF(struct RoseEngine, rolesWithStateCount); OF(struct RoseEngine, stateSize);
    OF(struct RoseEngine, anchorStateSize); OF(struct RoseEngine, tStateSize);
    OF(struct RoseEngine, scratchStateSize); OF(struct RoseEngine, smallWriteOffset);
    OF(struct RoseEngine, amatcherOffset); OF(struct RoseEngine, ematcherOffset);
    OF(struct RoseEngine, fmatcherOffset); OF(struct RoseEngine, drmatcherOffset);
    OF(struct RoseEngine, sbmatcherOffset); OF(struct RoseEngine, longLitTableOffset);
    OF(struct RoseEngine, amatcherMinWidth); OF(struct RoseEngine, fmatcherMinWidth);
    OF(struct RoseEngine, eodmatcherMinWidth); OF(struct RoseEngine, amatcherMaxBiAnchoredWidth);
    OF(struct RoseEngine, fmatcherMaxBiAnchoredWidth);
    OF(struct RoseEngine, reportProgramOffset); OF(struct RoseEngine, reportProgramCount);
    OF(struct RoseEngine, delayProgramOffset); OF(struct RoseEngine, anchoredProgramOffset);
    OF(struct RoseEngine, activeArrayCount); OF(struct RoseEngine, activeLeftCount);
    OF(struct RoseEngine, queueCount); OF(struct RoseEngine, activeQueueArraySize);
    OF(struct RoseEngine, eagerIterOffset); OF(struct RoseEngine, handledKeyCount);
    OF(struct RoseEngine, handledKeyFatbitSize); OF(struct RoseEngine, leftOffset);
    OF(struct RoseEngine, roseCount); OF(struct RoseEngine, eodProgramOffset);
    OF(struct RoseEngine, flushCombProgramOffset); OF(struct RoseEngine, lastFlushCombProgramOffset);
    OF(struct RoseEngine, lastByteHistoryIterOffset); OF(struct RoseEngine, minWidth);
    OF(struct RoseEngine, minWidthExcludingBoundaries); OF(struct RoseEngine, maxBiAnchoredWidth);
    OF(struct RoseEngine, anchoredDistance); OF(struct RoseEngine, anchoredMinDistance);
    OF(struct RoseEngine, floatingDistance); OF(struct RoseEngine, floatingMinDistance);
    OF(struct RoseEngine, smallBlockDistance); OF(struct RoseEngine, floatingMinLiteralMatchOffset);
    OF(struct RoseEngine, nfaInfoOffset); OF(struct RoseEngine, initialGroups);
    OF(struct RoseEngine, floating_group_mask); OF(struct RoseEngine, size);
    OF(struct RoseEngine, delay_count); OF(struct RoseEngine, delay_fatbit_size);
    OF(struct RoseEngine, anchored_count); OF(struct RoseEngine, anchored_fatbit_size);
    OF(struct RoseEngine, maxFloatingDelayedMatch); OF(struct RoseEngine, delayRebuildLength);
    OF(struct RoseEngine, stateOffsets); OF(struct RoseEngine, boundary);
    OF(struct RoseEngine, totalNumLiterals); OF(struct RoseEngine, asize);
    OF(struct RoseEngine, outfixBeginQueue); OF(struct RoseEngine, outfixEndQueue);
    OF(struct RoseEngine, leftfixBeginQueue); OF(struct RoseEngine, initMpvNfa);
    OF(struct RoseEngine, rosePrefixCount); OF(struct RoseEngine, activeLeftIterOffset);
    OF(struct RoseEngine, ematcherRegionSize); OF(struct RoseEngine, somRevCount);
    OF(struct RoseEngine, somRevOffsetOffset); OF(struct RoseEngine, longLitStreamState);
    OF(struct RoseEngine, state_init);
    SZ(struct RoseStateOffsets);
    OF(struct RoseStateOffsets, history); OF(struct RoseStateOffsets, exhausted);
    OF(struct RoseStateOffsets, exhausted_size); OF(struct RoseStateOffsets, logicalVec);
    OF(struct RoseStateOffsets, combVec); OF(struct RoseStateOffsets, activeLeafArray);
    OF(struct RoseStateOffsets, activeLeftArray); OF(struct RoseStateOffsets, leftfixLagTable);
    OF(struct RoseStateOffsets, anchorState); OF(struct RoseStateOffsets, groups);
    OF(struct RoseStateOffsets, groups_size); OF(struct RoseStateOffsets, longLitState);
    OF(struct RoseStateOffsets, somLocation); OF(struct RoseStateOffsets, somValid);
    OF(struct RoseStateOffsets, somWritable); OF(struct RoseStateOffsets, somMultibit_size);
    OF(struct RoseStateOffsets, nfaStateBegin); OF(struct RoseStateOffsets, end);
    SZ(struct RoseBoundaryReports);
    OF(struct RoseBoundaryReports, reportEodOffset); OF(struct RoseBoundaryReports, reportZeroOffset);
    OF(struct RoseBoundaryReports, reportZeroEodOffset);
    SZ(struct scatter_full_plan);
    VAL(ROSE_RUNTIME_FULL_ROSE); VAL(ROSE_RUNTIME_PURE_LITERAL); VAL(ROSE_RUNTIME_SINGLE_OUTFIX);
    VAL(ROSE_BOUND_INF);

    SZ(struct HWLM);
    OF(struct HWLM, type); OF(struct HWLM, accel1_groups); OF(struct HWLM, accel1); OF(struct HWLM, accel0);
    VAL(HWLM_ENGINE_NOOD); VAL(HWLM_ENGINE_FDR);
    SZ(union AccelAux);
    VAL(ACCEL_NONE); VAL(ACCEL_VERM); VAL(ACCEL_VERM_NOCASE); VAL(ACCEL_DVERM); VAL(ACCEL_DVERM_NOCASE);
    VAL(ACCEL_SHUFTI); VAL(ACCEL_DSHUFTI); VAL(ACCEL_TRUFFLE); VAL(ACCEL_RED_TAPE); VAL(ACCEL_DVERM_MASKED);
    SZ(struct noodTable);
    OF(struct noodTable, id); OF(struct noodTable, msk); OF(struct noodTable, cmp);
    OF(struct noodTable, msk_len); OF(struct noodTable, key_offset); OF(struct noodTable, nocase);
    OF(struct noodTable, single); OF(struct noodTable, key0); OF(struct noodTable, key1);

    SZ(struct FDR);
    OF(struct FDR, engineID); OF(struct FDR, size); OF(struct FDR, maxStringLen);
    OF(struct FDR, numStrings); OF(struct FDR, confOffset); OF(struct FDR, floodOffset);
    OF(struct FDR, stride); OF(struct FDR, domain); OF(struct FDR, domainMask);
    OF(struct FDR, tabSize); OF(struct FDR, start);
    SZ(struct FDRFlood);
    OF(struct FDRFlood, allGroups); OF(struct FDRFlood, suffix); OF(struct FDRFlood, idCount);
    OF(struct FDRFlood, ids); OF(struct FDRFlood, groups);
    SZ(struct FDRConfirm);
    OF(struct FDRConfirm, andmsk); OF(struct FDRConfirm, mult); OF(struct FDRConfirm, nBits);
    OF(struct FDRConfirm, groups);
    SZ(struct LitInfo);
    OF(struct LitInfo, v); OF(struct LitInfo, msk); OF(struct LitInfo, groups); OF(struct LitInfo, id);
    OF(struct LitInfo, size); OF(struct LitInfo, flags); OF(struct LitInfo, next);
    SZ(struct Teddy);
    OF(struct Teddy, engineID); OF(struct Teddy, size); OF(struct Teddy, maxStringLen);
    OF(struct Teddy, numStrings); OF(struct Teddy, confOffset); OF(struct Teddy, floodOffset);

    SZ(struct NFA);
    OF(struct NFA, flags); OF(struct NFA, length); OF(struct NFA, type); OF(struct NFA, rAccelType);
    OF(struct NFA, rAccelOffset); OF(struct NFA, maxBiAnchoredWidth); OF(struct NFA, rAccelData);
    OF(struct NFA, queueIndex); OF(struct NFA, nPositions); OF(struct NFA, scratchStateSize);
    OF(struct NFA, streamStateSize); OF(struct NFA, maxWidth); OF(struct NFA, minWidth);
    OF(struct NFA, maxOffset);
    VAL(LIMEX_NFA_32); VAL(LIMEX_NFA_64); VAL(LIMEX_NFA_128); VAL(LIMEX_NFA_256);
    VAL(LIMEX_NFA_384); VAL(LIMEX_NFA_512);

    SZ(struct hs_scratch);
    VAL(SCRATCH_MAGIC);

    /* Rose program instruction layouts (src/rose/rose_program.h:218-724) */
#define IOF(nm, t, f) printf("op." nm "." #f " %zu\n", offsetof(struct t, f))
#define ISZ(nm, t) printf("op." nm ".SIZE %zu\n", (size_t)ROUNDUP_N(sizeof(struct t), ROSE_INSTR_MIN_ALIGN))
    IOF("check_groups", ROSE_STRUCT_CHECK_GROUPS, groups); ISZ("check_groups", ROSE_STRUCT_CHECK_GROUPS);
    IOF("check_mask", ROSE_STRUCT_CHECK_MASK, and_mask); IOF("check_mask", ROSE_STRUCT_CHECK_MASK, cmp_mask);
    IOF("check_mask", ROSE_STRUCT_CHECK_MASK, neg_mask); IOF("check_mask", ROSE_STRUCT_CHECK_MASK, offset);
    IOF("check_mask", ROSE_STRUCT_CHECK_MASK, fail_jump); ISZ("check_mask", ROSE_STRUCT_CHECK_MASK);
    IOF("check_mask32", ROSE_STRUCT_CHECK_MASK_32, and_mask); IOF("check_mask32", ROSE_STRUCT_CHECK_MASK_32, cmp_mask);
    IOF("check_mask32", ROSE_STRUCT_CHECK_MASK_32, neg_mask); IOF("check_mask32", ROSE_STRUCT_CHECK_MASK_32, offset);
    IOF("check_mask32", ROSE_STRUCT_CHECK_MASK_32, fail_jump); ISZ("check_mask32", ROSE_STRUCT_CHECK_MASK_32);
    IOF("check_byte", ROSE_STRUCT_CHECK_BYTE, and_mask); IOF("check_byte", ROSE_STRUCT_CHECK_BYTE, cmp_mask);
    IOF("check_byte", ROSE_STRUCT_CHECK_BYTE, negation); IOF("check_byte", ROSE_STRUCT_CHECK_BYTE, offset);
    IOF("check_byte", ROSE_STRUCT_CHECK_BYTE, fail_jump); ISZ("check_byte", ROSE_STRUCT_CHECK_BYTE);
    IOF("check_shufti_16x8", ROSE_STRUCT_CHECK_SHUFTI_16x8, nib_mask); IOF("check_shufti_16x8", ROSE_STRUCT_CHECK_SHUFTI_16x8, bucket_select_mask);
    IOF("check_shufti_16x8", ROSE_STRUCT_CHECK_SHUFTI_16x8, neg_mask); IOF("check_shufti_16x8", ROSE_STRUCT_CHECK_SHUFTI_16x8, offset);
    IOF("check_shufti_16x8", ROSE_STRUCT_CHECK_SHUFTI_16x8, fail_jump); ISZ("check_shufti_16x8", ROSE_STRUCT_CHECK_SHUFTI_16x8);
    IOF("check_shufti_32x8", ROSE_STRUCT_CHECK_SHUFTI_32x8, hi_mask); IOF("check_shufti_32x8", ROSE_STRUCT_CHECK_SHUFTI_32x8, lo_mask);
    IOF("check_shufti_32x8", ROSE_STRUCT_CHECK_SHUFTI_32x8, bucket_select_mask); IOF("check_shufti_32x8", ROSE_STRUCT_CHECK_SHUFTI_32x8, neg_mask);
    IOF("check_shufti_32x8", ROSE_STRUCT_CHECK_SHUFTI_32x8, offset); IOF("check_shufti_32x8", ROSE_STRUCT_CHECK_SHUFTI_32x8, fail_jump);
    ISZ("check_shufti_32x8", ROSE_STRUCT_CHECK_SHUFTI_32x8);
    IOF("check_shufti_16x16", ROSE_STRUCT_CHECK_SHUFTI_16x16, hi_mask); IOF("check_shufti_16x16", ROSE_STRUCT_CHECK_SHUFTI_16x16, lo_mask);
    IOF("check_shufti_16x16", ROSE_STRUCT_CHECK_SHUFTI_16x16, bucket_select_mask); IOF("check_shufti_16x16", ROSE_STRUCT_CHECK_SHUFTI_16x16, neg_mask);
    IOF("check_shufti_16x16", ROSE_STRUCT_CHECK_SHUFTI_16x16, offset); IOF("check_shufti_16x16", ROSE_STRUCT_CHECK_SHUFTI_16x16, fail_jump);
    ISZ("check_shufti_16x16", ROSE_STRUCT_CHECK_SHUFTI_16x16);
    IOF("check_shufti_32x16", ROSE_STRUCT_CHECK_SHUFTI_32x16, hi_mask); IOF("check_shufti_32x16", ROSE_STRUCT_CHECK_SHUFTI_32x16, lo_mask);
    IOF("check_shufti_32x16", ROSE_STRUCT_CHECK_SHUFTI_32x16, bucket_select_mask_hi); IOF("check_shufti_32x16", ROSE_STRUCT_CHECK_SHUFTI_32x16, bucket_select_mask_lo);
    IOF("check_shufti_32x16", ROSE_STRUCT_CHECK_SHUFTI_32x16, neg_mask); IOF("check_shufti_32x16", ROSE_STRUCT_CHECK_SHUFTI_32x16, offset);
    IOF("check_shufti_32x16", ROSE_STRUCT_CHECK_SHUFTI_32x16, fail_jump); ISZ("check_shufti_32x16", ROSE_STRUCT_CHECK_SHUFTI_32x16);
    IOF("check_mask64", ROSE_STRUCT_CHECK_MASK_64, and_mask); IOF("check_mask64", ROSE_STRUCT_CHECK_MASK_64, cmp_mask);
    IOF("check_mask64", ROSE_STRUCT_CHECK_MASK_64, neg_mask); IOF("check_mask64", ROSE_STRUCT_CHECK_MASK_64, offset);
    IOF("check_mask64", ROSE_STRUCT_CHECK_MASK_64, fail_jump); ISZ("check_mask64", ROSE_STRUCT_CHECK_MASK_64);
    IOF("check_shufti_64x8", ROSE_STRUCT_CHECK_SHUFTI_64x8, hi_mask); IOF("check_shufti_64x8", ROSE_STRUCT_CHECK_SHUFTI_64x8, lo_mask);
    IOF("check_shufti_64x8", ROSE_STRUCT_CHECK_SHUFTI_64x8, bucket_select_mask); IOF("check_shufti_64x8", ROSE_STRUCT_CHECK_SHUFTI_64x8, neg_mask);
    IOF("check_shufti_64x8", ROSE_STRUCT_CHECK_SHUFTI_64x8, offset); IOF("check_shufti_64x8", ROSE_STRUCT_CHECK_SHUFTI_64x8, fail_jump);
    ISZ("check_shufti_64x8", ROSE_STRUCT_CHECK_SHUFTI_64x8);
    IOF("check_shufti_64x16", ROSE_STRUCT_CHECK_SHUFTI_64x16, hi_mask_1); IOF("check_shufti_64x16", ROSE_STRUCT_CHECK_SHUFTI_64x16, hi_mask_2);
    IOF("check_shufti_64x16", ROSE_STRUCT_CHECK_SHUFTI_64x16, lo_mask_1); IOF("check_shufti_64x16", ROSE_STRUCT_CHECK_SHUFTI_64x16, lo_mask_2);
    IOF("check_shufti_64x16", ROSE_STRUCT_CHECK_SHUFTI_64x16, bucket_select_mask_hi); IOF("check_shufti_64x16", ROSE_STRUCT_CHECK_SHUFTI_64x16, bucket_select_mask_lo);
    IOF("check_shufti_64x16", ROSE_STRUCT_CHECK_SHUFTI_64x16, neg_mask); IOF("check_shufti_64x16", ROSE_STRUCT_CHECK_SHUFTI_64x16, offset);
    IOF("check_shufti_64x16", ROSE_STRUCT_CHECK_SHUFTI_64x16, fail_jump); ISZ("check_shufti_64x16", ROSE_STRUCT_CHECK_SHUFTI_64x16);
    IOF("dedupe", ROSE_STRUCT_DEDUPE, quash_som); IOF("dedupe", ROSE_STRUCT_DEDUPE, dkey);
    IOF("dedupe", ROSE_STRUCT_DEDUPE, offset_adjust); IOF("dedupe", ROSE_STRUCT_DEDUPE, fail_jump); ISZ("dedupe", ROSE_STRUCT_DEDUPE);
    IOF("report", ROSE_STRUCT_REPORT, onmatch); IOF("report", ROSE_STRUCT_REPORT, offset_adjust); ISZ("report", ROSE_STRUCT_REPORT);
    IOF("final_report", ROSE_STRUCT_FINAL_REPORT, onmatch); IOF("final_report", ROSE_STRUCT_FINAL_REPORT, offset_adjust); ISZ("final_report", ROSE_STRUCT_FINAL_REPORT);
    IOF("report_exhaust", ROSE_STRUCT_REPORT_EXHAUST, onmatch); IOF("report_exhaust", ROSE_STRUCT_REPORT_EXHAUST, offset_adjust);
    IOF("report_exhaust", ROSE_STRUCT_REPORT_EXHAUST, ekey); ISZ("report_exhaust", ROSE_STRUCT_REPORT_EXHAUST);
    IOF("dedupe_and_report", ROSE_STRUCT_DEDUPE_AND_REPORT, quash_som); IOF("dedupe_and_report", ROSE_STRUCT_DEDUPE_AND_REPORT, dkey);
    IOF("dedupe_and_report", ROSE_STRUCT_DEDUPE_AND_REPORT, onmatch); IOF("dedupe_and_report", ROSE_STRUCT_DEDUPE_AND_REPORT, offset_adjust);
    IOF("dedupe_and_report", ROSE_STRUCT_DEDUPE_AND_REPORT, fail_jump); ISZ("dedupe_and_report", ROSE_STRUCT_DEDUPE_AND_REPORT);
    IOF("check_exhausted", ROSE_STRUCT_CHECK_EXHAUSTED, ekey); IOF("check_exhausted", ROSE_STRUCT_CHECK_EXHAUSTED, fail_jump);
    ISZ("check_exhausted", ROSE_STRUCT_CHECK_EXHAUSTED);
    IOF("squash_groups", ROSE_STRUCT_SQUASH_GROUPS, groups); ISZ("squash_groups", ROSE_STRUCT_SQUASH_GROUPS);
    IOF("check_lit", ROSE_STRUCT_CHECK_MED_LIT, lit_offset); IOF("check_lit", ROSE_STRUCT_CHECK_MED_LIT, lit_length);
    IOF("check_lit", ROSE_STRUCT_CHECK_MED_LIT, fail_jump); ISZ("check_lit", ROSE_STRUCT_CHECK_MED_LIT);
    IOF("included_jump", ROSE_STRUCT_INCLUDED_JUMP, squash); IOF("included_jump", ROSE_STRUCT_INCLUDED_JUMP, child_offset);
    ISZ("included_jump", ROSE_STRUCT_INCLUDED_JUMP);
    IOF("set_exhaust", ROSE_STRUCT_SET_EXHAUST, ekey); ISZ("set_exhaust", ROSE_STRUCT_SET_EXHAUST);
    ISZ("end", ROSE_STRUCT_END);
    VAL(ROSE_INSTR_END); VAL(ROSE_INSTR_CHECK_GROUPS); VAL(ROSE_INSTR_CHECK_MASK); VAL(ROSE_INSTR_CHECK_MASK_32);
    VAL(ROSE_INSTR_CHECK_BYTE); VAL(ROSE_INSTR_CHECK_SHUFTI_16x8); VAL(ROSE_INSTR_DEDUPE); VAL(ROSE_INSTR_REPORT);
    VAL(ROSE_INSTR_REPORT_EXHAUST); VAL(ROSE_INSTR_DEDUPE_AND_REPORT); VAL(ROSE_INSTR_FINAL_REPORT);
    VAL(ROSE_INSTR_CHECK_EXHAUSTED); VAL(ROSE_INSTR_SQUASH_GROUPS); VAL(ROSE_INSTR_CHECK_LONG_LIT);
    VAL(ROSE_INSTR_CHECK_MED_LIT); VAL(ROSE_INSTR_CHECK_MED_LIT_NOCASE); VAL(ROSE_INSTR_CLEAR_WORK_DONE);
    VAL(ROSE_INSTR_INCLUDED_JUMP); VAL(ROSE_INSTR_SET_EXHAUST); VAL(ROSE_INSTR_CHECK_SHUFTI_64x8);
    VAL(ROSE_INSTR_CHECK_MASK_64); VAL(LAST_ROSE_INSTRUCTION);
    /* AccelAux variants (src/nfa/accel.h:72-142) */
    printf("accel.verm_c %zu\n", offsetof(union AccelAux, verm.c));
    printf("accel.dverm_c1 %zu\n", offsetof(union AccelAux, dverm.c1));
    printf("accel.dverm_c2 %zu\n", offsetof(union AccelAux, dverm.c2));
    printf("accel.dverm_m1 %zu\n", offsetof(union AccelAux, dverm.m1));
    printf("accel.dverm_m2 %zu\n", offsetof(union AccelAux, dverm.m2));
    printf("accel.shufti_lo %zu\n", offsetof(union AccelAux, shufti.lo));
    printf("accel.shufti_hi %zu\n", offsetof(union AccelAux, shufti.hi));
    printf("accel.truffle_mask1 %zu\n", offsetof(union AccelAux, truffle.mask1));
    printf("accel.truffle_mask2 %zu\n", offsetof(union AccelAux, truffle.mask2));
    printf("accel.offset %zu\n", offsetof(union AccelAux, generic.offset));
    OF(struct RoseStateOffsets, logicalVec_size); OF(struct RoseStateOffsets, combVec_size);
    OF(struct RoseStateOffsets, activeLeafArray_size); OF(struct RoseStateOffsets, activeLeftArray_size);
    OF(struct RoseStateOffsets, longLitState_size);
    OF(struct RoseEngine, logicalTreeOffset); OF(struct RoseEngine, combInfoMapOffset);
    OF(struct hs_database, reserved0); OF(struct hs_database, reserved1); OF(struct hs_database, padding);
    VAL(HS_MODE_BLOCK); VAL(HS_MODE_STREAM); VAL(HS_MODE_VECTORED);
    VAL(FDR_FLOOD_MAX_IDS); VAL(FDR_LIT_FLAG_NOREPEAT);
    /* NFA engines */

 SZ(struct NFA); OF(struct NFA, flags); OF(struct NFA, length); OF(struct NFA, type); OF(struct NFA, rAccelType); OF(struct NFA, rAccelOffset);
 OF(struct NFA, maxBiAnchoredWidth); OF(struct NFA, rAccelData); OF(struct NFA, queueIndex); OF(struct NFA, nPositions); OF(struct NFA, scratchStateSize);
 OF(struct NFA, streamStateSize); OF(struct NFA, maxWidth); OF(struct NFA, minWidth); OF(struct NFA, maxOffset);
 VAL(LIMEX_NFA_32); VAL(LIMEX_NFA_64); VAL(LIMEX_NFA_128); VAL(LIMEX_NFA_256); VAL(LIMEX_NFA_384); VAL(LIMEX_NFA_512);
 VAL(NFA_ACCEPTS_EOD); VAL(NFA_ZOMBIE); VAL(MO_INVALID_IDX);
 LIMEX(32) LIMEX(64) LIMEX(128) LIMEX(256) LIMEX(384) LIMEX(512)
 SZ(struct NFAAccept); OF(struct NFAAccept, single_report); OF(struct NFAAccept, reports); OF(struct NFAAccept, squash);
 SZ(struct NfaInfo); OF(struct NfaInfo, nfaOffset); OF(struct NfaInfo, stateOffset); OF(struct NfaInfo, fullStateOffset); 
 OF(struct NfaInfo, no_retrigger); OF(struct NfaInfo, in_sbmatcher); OF(struct NfaInfo, eod);
 /* Sheng DFA */
 VAL(SHENG_NFA); VAL(SHENG_STATE_ACCEPT); VAL(SHENG_STATE_DEAD); VAL(SHENG_STATE_ACCEL); VAL(SHENG_STATE_MASK);
 VAL(SHENG_FLAG_SINGLE_REPORT); VAL(SHENG_FLAG_CAN_DIE); VAL(SHENG_FLAG_HAS_ACCEL);
 SZ(struct sheng); OF(struct sheng, shuffle_masks); OF(struct sheng, length); OF(struct sheng, aux_offset);
 OF(struct sheng, report_offset); OF(struct sheng, accel_offset); OF(struct sheng, n_states); OF(struct sheng, anchored);
 OF(struct sheng, floating); OF(struct sheng, flags); OF(struct sheng, report);
 SZ(struct sstate_aux); OF(struct sstate_aux, accept); OF(struct sstate_aux, accept_eod); OF(struct sstate_aux, accel);
 OF(struct sstate_aux, top);
 /* McClellan DFA */
 VAL(MCCLELLAN_NFA_8); VAL(MCCLELLAN_NFA_16); VAL(ACCEPT_FLAG); VAL(ACCEL_FLAG); VAL(STATE_MASK); VAL(MCCLELLAN_FLAG_SINGLE);
 VAL(SHERMAN_STATE); VAL(SHERMAN_FIXED_SIZE); VAL(SHERMAN_LEN_OFFSET); VAL(SHERMAN_DADDY_OFFSET); VAL(SHERMAN_CHARS_OFFSET);
 SZ(struct mcsheng); OF(struct mcsheng, state_count); OF(struct mcsheng, length); OF(struct mcsheng, start_anchored);
 OF(struct mcsheng, start_floating); OF(struct mcsheng, aux_offset); OF(struct mcsheng, sherman_offset);
 OF(struct mcsheng, sherman_end); OF(struct mcsheng, sheng_end); OF(struct mcsheng, sheng_accel_limit);
 OF(struct mcsheng, accel_limit_8); OF(struct mcsheng, accept_limit_8); OF(struct mcsheng, sherman_limit);
 OF(struct mcsheng, alphaShift); OF(struct mcsheng, flags); OF(struct mcsheng, has_accel); OF(struct mcsheng, remap);
 OF(struct mcsheng, arb_report); OF(struct mcsheng, accel_offset); OF(struct mcsheng, sheng_masks);
 VAL(MCSHENG_NFA_8); VAL(MCSHENG_NFA_16); VAL(MCSHENG_64_NFA_8); VAL(MCSHENG_64_NFA_16);
 SZ(struct mcsheng64); OF(struct mcsheng64, sheng_end); OF(struct mcsheng64, remap); OF(struct mcsheng64, arb_report);
 OF(struct mcsheng64, accel_offset); OF(struct mcsheng64, sheng_succ_masks);
 SZ(struct mcclellan); OF(struct mcclellan, state_count); OF(struct mcclellan, length); OF(struct mcclellan, start_anchored);
 OF(struct mcclellan, start_floating); OF(struct mcclellan, aux_offset); OF(struct mcclellan, sherman_offset);
 OF(struct mcclellan, sherman_end); OF(struct mcclellan, accel_limit_8); OF(struct mcclellan, accept_limit_8);
 OF(struct mcclellan, sherman_limit); OF(struct mcclellan, wide_limit); OF(struct mcclellan, alphaShift);
 OF(struct mcclellan, flags); OF(struct mcclellan, has_accel); OF(struct mcclellan, has_wide); OF(struct mcclellan, remap);
 OF(struct mcclellan, arb_report); OF(struct mcclellan, accel_offset); OF(struct mcclellan, haig_offset);
 OF(struct mcclellan, wide_offset);
 SZ(struct mstate_aux); OF(struct mstate_aux, accept); OF(struct mstate_aux, accept_eod); OF(struct mstate_aux, top);
 OF(struct mstate_aux, accel_offset);


    return 0;
}
