"""Summarise an .ncu-rep here (no GPU): per-launch key metrics + per-opcode issue histogram.
    python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [kernel-index-for-opcode-histogram]"""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keys = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__registers_per_thread', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__inst_executed.sum', 'smsp__thread_inst_executed.sum',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio']
ki = hdr.index('Kernel Name')
for n, r in enumerate(rows[2:]):
    print("== launch %d: %s" % (n, r[ki][:70]))
    for k in keys:
        if k in hdr:
            i = hdr.index(k)
            print("   %-85s %14s %s" % (k, r[i], units[i]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks, cur = [], None
for r in csv.reader(io.StringIO(src)):
    if r and r[0] == 'Kernel Name': cur = {'name': r[1], 'rows': []}; blocks.append(cur); continue
    if r and r[0] == 'Address': cur['hdr'] = r; continue
    if cur is not None and len(r) > 10: cur['rows'].append(r)
# ncu prints each kernel twice (SASS view + source view); keep those with SASS rows
blocks = [b for b in blocks if b['rows'] and b['rows'][0][0].startswith('0x')]
seen = []
for b in blocks:
    if any(b['rows'][0][0] == s['rows'][0][0] and b['name'] == s['name'] and len(b['rows']) == len(s['rows']) and b['rows'][0][5] == s['rows'][0][5] for s in seen): continue
    seen.append(b)
which = int(sys.argv[2]) if len(sys.argv) > 2 else None
for n, b in enumerate(seen):
    h = b['hdr']; ie = h.index('Instructions Executed'); te = h.index('Thread Instructions Executed'); s = h.index('Source')
    tot = sum(int(r[ie]) for r in b['rows']); tt = sum(int(r[te]) for r in b['rows'])
    print("== sass %d: %s  warp-insts %d thread-insts %d lanes/inst %.2f  sass lines %d" % (n, b['name'][:60], tot, tt, tt / max(1, tot), len(b['rows'])))
    if which is None or which == n:
        op = collections.Counter()
        for r in b['rows']:
            o = [x for x in r[s].split() if not x.startswith('@')][0].split('.')[0]
            op[o] += int(r[ie])
        print("   " + "  ".join("%s %.1f%%" % (k, 100 * v / tot) for k, v in op.most_common(16)))
