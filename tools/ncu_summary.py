"""Summarise an .ncu-rep (first kernel): python tools/ncu_summary.py <rep> [--segments]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, d = rows[0], rows[1], dict(zip(rows[0], rows[2]))
want = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "sm__cycles_elapsed.max", "smsp__cycles_active.avg", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sectors_srcunit_tex_op_read.sum",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts_mem_lgds.avg",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"]
for k in want:
    if k in d:
        print(f"{k}: {d[k]} {units[hdr.index(k)]}")
for k in hdr:
    if "issue_stalled" in k and "per_issue_active" in k:
        v = float(d[k] or 0)
        if v >= 0.3:
            print(f"{k.replace('smsp__average_warps_issue_stalled_', 'stall ').replace('_per_issue_active.ratio', '')}: {v:.2f}")
if "--segments" in sys.argv:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}; data = rows[2:]
    tot = sum(int(r[ix['Instructions Executed']]) for r in data); tots = sum(int(r[ix['# Samples']]) for r in data)
    print("instructions", tot, "samples", tots)
    segs = []; cur = None
    for i, r in enumerate(data):
        n = int(r[ix['Instructions Executed']]); s = int(r[ix['# Samples']])
        wf = int(r[ix['L1 Wavefronts Shared']] or 0)
        if cur and abs(n - cur['ref']) <= 0.15 * max(cur['ref'], 1):
            cur['n'] += n; cur['s'] += s; cur['end'] = i; cur['wf'] += wf
        else:
            cur = {'start': i, 'end': i, 'ref': n, 'n': n, 's': s, 'wf': wf}; segs.append(cur)
    for sg in segs:
        if sg['n'] > 0.004 * tot or sg['s'] > 0.006 * tots:
            print(f"{sg['start']:5d}-{sg['end']:5d} len {sg['end']-sg['start']+1:4d} x{sg['ref']//1000:6d}k inst% {100*sg['n']/tot:5.1f} samp% {100*sg['s']/tots:5.1f} smem_wf {sg['wf']//1000:6d}k  {data[sg['start']][ix['Source']][:44]}")
