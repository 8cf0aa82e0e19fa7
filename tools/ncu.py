"""Summarise an ncu report: per-launch headline metrics and (optionally) per-region stall samples."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_uniform.sum',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__inst_executed.sum', 'launch__registers_per_thread', 'launch__grid_size',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__warp_issue_stalled_no_instruction_per_warp_active.pct']
for w in want:
    if w in idx:
        print("%-66s %-7s" % (w[:66], units[idx[w]]), [r[idx[w]][:28] for r in rows[2:]])
if len(sys.argv) > 2:
    kid = sys.argv[2]
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", kid], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hdr = rows[1]; idx = {h: i for i, h in enumerate(hdr)}
    data = []
    for r in rows[2:]:
        if r and r[0] == 'Kernel Name': break
        if len(r) == len(hdr): data.append(r)
    stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
    B = int(sys.argv[3]) if len(sys.argv) > 3 else 100
    tot = sum(int(r[idx['# Samples']] or 0) for r in data)
    print("kernel", rows[0][1][:80], "instrs", len(data), "samples", tot)
    cur = None
    for i, r in enumerate(data):
        if i % B == 0:
            if cur: print(cur['start'], cur['samples'], cur['exec'], sorted(cur['marks']), sorted(cur['st'].items(), key=lambda x: -x[1])[:3])
            cur = {'samples': 0, 'marks': set(), 'start': i, 'st': {}, 'exec': 0}
        s = int(r[idx['# Samples']] or 0)
        cur['samples'] += s
        cur['exec'] = max(cur['exec'], int(r[idx['Instructions Executed']] or 0))
        for st in stalls:
            v = int(r[idx[st]] or 0)
            if v: cur['st'][st[6:]] = cur['st'].get(st[6:], 0) + v
        for m in ('UTCHMMA', 'LDTM', 'LDGSTS', 'UBLKCP', 'UTCBAR', 'STG', 'BAR.SYNC', 'EXIT', 'TRYWAIT', 'ARRIVE', 'R2UR'):
            if m in r[idx['Source']]: cur['marks'].add(m)
    print(cur['start'], cur['samples'], cur['exec'], sorted(cur['marks']), sorted(cur['st'].items(), key=lambda x: -x[1])[:3])

# every mbarrier wait site (first TRYWAIT of each mbar_wait) with the samples of its poll loop
if len(sys.argv) > 2:
    print("wait sites (instr index, samples in [site, site+32), executions of first try_wait):")
    i = 0
    while i < len(data):
        if 'TRYWAIT' in data[i][idx['Source']]:
            smp = sum(int(r[idx['# Samples']] or 0) for r in data[i:i + 32])
            print("  %5d  samples %6d  exec %s" % (i, smp, data[i][idx['Instructions Executed']]))
            i += 32
        else:
            i += 1
