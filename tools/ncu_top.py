"""Print the hottest SASS instructions (stall samples) of each kernel instance in an ncu report (dev tool).
    python tools/ncu_top.py report.ncu-rep [N]"""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 22
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
if len(rows) > 2:
    hdr, units = rows[0], rows[1]
    want = ['gpu__time_duration.sum', 'sm__cycles_elapsed.max', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct',
            'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
            'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
            'launch__registers_per_thread', 'smsp__inst_executed.sum', 'l1tex__data_pipe_lsu_wavefronts.sum',
            'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active']
    for i, h in enumerate(hdr):
        if h in want:
            print(f"{h} [{units[i]}]: " + " | ".join(r[i][:14] for r in rows[2:]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
inst = []; cur = None
for r in rows:
    if r and r[0] == 'Kernel Name':
        cur = {'name': r[1][:60], 'hdr': None, 'rows': []}; inst.append(cur); continue
    if cur is None: continue
    if cur['hdr'] is None: cur['hdr'] = r; continue
    cur['rows'].append(r)
for k in inst:
    h, R = k['hdr'], k['rows']
    iS, iSrc, iE = h.index('# Samples'), h.index('Source'), h.index('Instructions Executed')
    sc = [i for i, c in enumerate(h) if c.startswith('stall_') and 'Not Issued' not in c]
    tot = sum(int(r[iS]) for r in R); ex = sum(int(r[iE]) for r in R)
    print(f"== {k['name']}  samples={tot} warp-instr={ex}")
    agg = collections.Counter()
    for r in R:
        for i in sc: agg[h[i]] += int(r[i])
    print("   stall totals:", ", ".join(f"{a}={b}" for a, b in agg.most_common(8)))
    for n, r in sorted(enumerate(R), key=lambda t: -int(t[1][iS]))[:N]:
        st = sorted(((h[i], int(r[i])) for i in sc if int(r[i]) > 0), key=lambda kv: -kv[1])[:2]
        print(f"  {n:5d} {int(r[iS]):6d} {100*int(r[iS])/max(tot,1):5.1f}% ex={r[iE]:>8s} {r[iSrc].strip()[:58]:58s} {st}")
