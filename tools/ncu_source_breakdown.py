"""Per-file / per-line breakdown of an .ncu-rep source page: executed warp instructions, lane efficiency, stall samples without
the block barrier, memory spaces.   python tools/ncu_source_breakdown.py gpurun_out/env_step.ncu-rep [topN]"""
import collections, csv, subprocess, sys


def main(rep, topn=40):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows = csv.reader(out.splitlines())
    cur = None; hdr = None
    F = collections.defaultdict(lambda: collections.Counter())
    Lines = collections.defaultdict(lambda: collections.Counter()); src = {}
    space = collections.Counter()
    curline = None
    for r in rows:
        if not r: continue
        if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
        if r[0] == 'Function Name': continue
        if r[0] == 'Line No':
            hdr = r
            ix = {n: i for i, n in enumerate(hdr)}
            stalls = [(i, x) for i, x in enumerate(hdr) if x.startswith('stall_') and 'Not Issued' not in x]
            continue
        def f(name):
            try: return float(r[ix[name]] or 0)
            except (ValueError, IndexError): return 0.0
        if r[0] == '':      # SASS row under the current CUDA line
            sp = r[ix['Address Space']]
            if sp:
                key = (sp, r[ix['Access Operation']])
                space[key] += f('Instructions Executed')
                if curline: Lines[curline]['mem_' + sp] += f('Instructions Executed')
            continue
        curline = (cur, r[0]); src[curline] = r[1].strip()[:110]
        c = Lines[curline]
        c['inst'] += f('Instructions Executed'); c['tinst'] += f('Thread Instructions Executed'); c['samples'] += f('# Samples')
        for i, x in stalls:
            try: v = float(r[i] or 0)
            except ValueError: v = 0
            c[x] += v
        for k, v in c.items(): pass
    for k, c in Lines.items():
        for kk, v in c.items(): F[k[0]][kk] += v
    tot = collections.Counter()
    for c in F.values():
        for kk, v in c.items(): tot[kk] += v
    nb = lambda c: c['samples'] - c['stall_barrier']
    print("total warp inst %.3g  thread inst %.3g  lanes/inst %.2f  samples %d (barrier %.1f %%)" % (tot['inst'], tot['tinst'], tot['tinst'] / max(tot['inst'], 1), tot['samples'], 100 * tot['stall_barrier'] / max(tot['samples'], 1)))
    print("\nper file: %inst  lanes/inst  %non-barrier samples   [long_sb short_sb wait no_inst branch lg] as % of the file's non-barrier samples")
    for fn, c in sorted(F.items(), key=lambda kv: -nb(kv[1])):
        n = max(nb(c), 1)
        print("  %-18s %5.1f %%  %5.1f   %5.1f %%   [%4.0f %4.0f %4.0f %4.0f %4.0f %4.0f]" % (fn, 100 * c['inst'] / tot['inst'], c['tinst'] / max(c['inst'], 1), 100 * nb(c) / nb(tot),
              100 * c['stall_long_sb'] / n, 100 * c['stall_short_sb'] / n, 100 * c['stall_wait'] / n, 100 * c['stall_no_inst'] / n, 100 * c['stall_branch_resolving'] / n, 100 * c['stall_lg'] / n))
    print("\nmemory instructions by space/op (warp instructions):")
    for k, v in space.most_common(12): print("  %-22s %.3g" % (k, v))
    print("\ntop lines by non-barrier samples: %samples  %inst  lanes  dominant stall")
    for k, c in sorted(Lines.items(), key=lambda kv: -nb(kv[1]))[:topn]:
        dom = max(((x, c[x]) for i, x in stalls if x != 'stall_barrier'), key=lambda t: t[1])
        print("  %5.2f %%  %5.2f %%  %4.1f  %-14s %s:%s  %s" % (100 * nb(c) / nb(tot), 100 * c['inst'] / tot['inst'], c['tinst'] / max(c['inst'], 1), dom[0][6:], k[0], k[1], src[k]))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
