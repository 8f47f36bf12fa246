#!/usr/bin/env python
"""Key figures + stall-sample hot spots of each kernel in an ncu report:  python scripts/ncu_kernel_report.py <report.ncu-rep> [kernel regex]"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]; pat = sys.argv[2] if len(sys.argv) > 2 else "."
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); hdr, units, data = rows[0], rows[1], rows[2:]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'smsp__inst_executed.sum',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'launch__waves_per_multiprocessor', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'smsp__cycles_active.avg', 'sm__cycles_elapsed.max']
kn = hdr.index('Kernel Name')
for r in data:
    if not re.search(pat, r[kn]): continue
    print("==", r[kn][:90])
    for w in want:
        if w in hdr: print("   %-62s %s %s" % (w, r[hdr.index(w)], units[hdr.index(w)]))
    for i, h in enumerate(hdr):
        if 'issue_stalled' in h and h.endswith('per_issue_active.ratio'):
            try:
                if float(r[i]) > 0.3: print("   stall %-40s %s" % (h.split('issue_stalled_')[1].split('_per_')[0], r[i]))
            except ValueError: pass
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
name = None; hdr = None; seen = set(); insts = []
def flush():
    if not insts or name in seen or not re.search(pat, name): return
    seen.add(name)
    iS, iSrc = hdr.index('# Samples'), hdr.index('Source'); st = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
    tot = sum(int(r[iS]) for r in insts)
    print("== samples by instruction:", name[:80], "total", tot)
    for i in sorted(sorted(range(len(insts)), key=lambda i: -int(insts[i][iS]))[:22]):
        r = insts[i]; why = max(st, key=lambda j: int(r[j] or 0))
        print("   %5d %-62s %5s  %s" % (i, r[iSrc].strip()[:62], r[iS], hdr[why]))
    for s in range(0, len(insts), 200): print("   [%d..) %d" % (s, sum(int(r[iS]) for r in insts[s:s + 200])))
for r in csv.reader(io.StringIO(src)):
    if r and r[0] == 'Kernel Name': flush(); name = r[1]; insts = []; hdr = None; continue
    if r and r[0] == 'Address': 
        if hdr is not None: flush(); insts = None
        hdr = r; 
        if insts is None: insts = []
        continue
    if hdr and insts is not None and len(r) >= len(hdr) - 2 and name not in seen: insts.append(r)
flush()
