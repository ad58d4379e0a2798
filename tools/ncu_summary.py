"""Text summary of an .ncu-rep (raw metrics of the first launch + whole-kernel stall breakdown + top source lines):
    python tools/ncu_summary.py gpurun_out/env_step.ncu-rep > profiles/<name>.txt"""
import collections, csv, subprocess, sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'smsp__inst_executed.sum', 'sm__inst_executed_pipe_fma.sum.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'launch__waves_per_multiprocessor',
        'smsp__warps_eligible.avg.per_cycle_active', 'sass__inst_executed_local_loads', 'sass__inst_executed_local_stores',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct']


def run(args):
    return subprocess.run(["ncu", "-i"] + args, capture_output=True, text=True).stdout


def main(rep):
    rows = list(csv.reader(run([rep, "--page", "raw", "--csv"]).splitlines()))
    h, u, v = rows[0], rows[1], rows[2]
    print("kernel:", v[h.index("Kernel Name")])
    for a, b, c in zip(h, u, v):
        if a in WANT:
            print("  %-72s %-18s %s" % (a, b, c))
    rows = list(csv.reader(run([rep, "--page", "source", "--csv", "--print-source", "cuda,sass"]).splitlines()))
    cur = None; hdr = None
    byline = collections.Counter(); src = {}; tot = collections.Counter(); nsass = 0
    for r in rows:
        if not r: continue
        if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
        if r[0] == 'Function Name': continue
        if r[0] == 'Line No':
            hdr = r; i_s = hdr.index('# Samples')
            st = [(i, x) for i, x in enumerate(hdr) if x.startswith('stall_') and 'Not Issued' not in x]
            continue
        if r[0] == '': nsass += 1; continue
        try: s = float(r[i_s] or 0)
        except ValueError: continue
        byline[(cur, r[0])] += s; src[(cur, r[0])] = r[1].strip()[:100]
        for i, x in st:
            try: tot[x] += float(r[i] or 0)
            except ValueError: pass
    S = sum(tot.values()) or 1.0; T = sum(byline.values()) or 1.0
    print("\nSASS instructions:", nsass)
    print("warp stall sampling, whole kernel:")
    for x, val in tot.most_common(9):
        print("  %-26s %5.1f %%" % (x, 100 * val / S))
    print("\nsource lines with the most samples:")
    for k, s in byline.most_common(14):
        print("  %5.2f %%  %s:%s  %s" % (100 * s / T, k[0], k[1], src[k]))


if __name__ == "__main__":
    main(sys.argv[1])
