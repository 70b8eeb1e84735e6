"""Helpers to summarise ncu exports (raw page CSV, source page CSV with --print-source cuda,sass)."""
import csv, sys

def raw_summary(path, want=None):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    want = want or ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
                    'smsp__issue_active.avg.pct', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
                    'launch__occupancy_limit', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
                    'smsp__average_warps_issue_stalled', 'l1tex__t_sector_hit_rate', 'lts__t_sector_hit_rate', 'sm__inst_executed.avg.per_cycle_active',
                    'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__inst_executed_op_shared', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum']
    for i, h in enumerate(hdr):
        if any(h.startswith(w) for w in want) and 'per_second' not in h and 'pct_of_peak' not in h.replace('sm__warps_active.avg.pct_of_peak', ''):
            vals = [d[i] for d in data]
            if any(v not in ('0', '0.000000') for v in vals):
                print(f"{h} [{units[i]}] {vals}")

def source_lines(path):
    rows = list(csv.reader(open(path)))
    cur = None; agg = {}
    for r in rows:
        if not r: continue
        if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
        if r[0] in ('Function Name', 'Line No'): continue
        if len(r) < 10: continue
        if r[2] == '-' and r[0].isdigit():
            agg[(cur, int(r[0]))] = (int(r[7]), int(r[8]), int(r[6]), r[1])
    return agg

def top_lines(path, n=40):
    agg = source_lines(path)
    tot = sum(v[0] for v in agg.values())
    print('total inst', tot)
    for (f, ln), (inst, tinst, samp, src) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:n]:
        print(f"{f}:{ln:4d} inst={inst/1e6:7.2f}M ({100*inst/tot:4.1f}%) eff={tinst/max(inst,1):5.1f} samp={samp:5d} | {src.strip()[:100]}")

def ranges(path, spec):
    agg = source_lines(path)
    tot = sum(v[0] for v in agg.values())
    for name, f, a, b in spec:
        i = sum(v[0] for (ff, l), v in agg.items() if ff == f and a <= l <= b)
        t = sum(v[1] for (ff, l), v in agg.items() if ff == f and a <= l <= b)
        s = sum(v[2] for (ff, l), v in agg.items() if ff == f and a <= l <= b)
        print(f"{name:12s} inst={i/1e6:7.1f}M {100*i/tot:5.1f}% eff={t/max(i,1):5.1f} samples={s}")

if __name__ == '__main__':
    if sys.argv[1] == 'raw': raw_summary(sys.argv[2])
    elif sys.argv[1] == 'top': top_lines(sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 40)


def sass_rows(path):
    """[(address, file, line, sass, warp_inst, thread_inst)] sorted by address, from a cuda,sass source-page CSV."""
    rows = list(csv.reader(open(path)))
    cur = None; line = None; out = {}
    for r in rows:
        if not r: continue
        if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
        if r[0] in ('Function Name', 'Line No') or len(r) < 10: continue
        if r[0].strip().isdigit() and r[2] == '-': line = int(r[0]); continue
        a = r[2].strip()
        if a.startswith('0x'):
            out[int(a, 16)] = (cur, line, r[3].strip(), int(r[7]), int(r[8]))
    return [(a,) + v for a, v in sorted(out.items())]


def dump_sass(path, f, l0, l1):
    rows = sass_rows(path)
    sel = [a for a, ff, l, *_ in rows if ff == f and l0 <= l <= l1]
    lo, hi = min(sel), max(sel)
    for a, ff, l, s, i, t in rows:
        if lo <= a <= hi:
            print(f"{a & 0xfffff:05x} {ff[:12]}:{l:3d} {i/1e3:8.0f}k eff={t/max(i,1):4.1f}  {s}")
