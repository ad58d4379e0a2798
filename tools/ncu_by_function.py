"""Dynamic warp instructions / lane efficiency / stall samples of an .ncu-rep source page aggregated by the FUNCTION each source
line belongs to (function = the nearest preceding GGL_HD / GGL_HDN / __global__ definition in the file).
    python tools/ncu_by_function.py gpurun_out/env_step.ncu-rep [topN]"""
import collections, csv, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "reinforcement-learning_b200", "csrc")


def func_map(fn):
    p = os.path.join(CSRC, fn)
    if not os.path.exists(p):
        return None
    starts = []
    for i, l in enumerate(open(p).read().splitlines(), 1):
        m = re.match(r"\s*(?:template\s*<[^>]*>\s*)?(?:GGL_HDN?|__global__|__device__|inline|static)\b.*?\b([A-Za-z_][A-Za-z0-9_]*)\s*\(", l)
        if m and not l.strip().startswith("//"):
            starts.append((i, m.group(1)))
    return starts


def main(rep, topn=45):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda"], capture_output=True, text=True).stdout
    rows = csv.reader(out.splitlines())
    cur = None; ix = None
    agg = collections.defaultdict(collections.Counter)
    maps = {}
    for r in rows:
        if not r: continue
        if r[0] == 'File Path': cur = os.path.basename(r[1]); continue
        if r[0] == 'Function Name': continue
        if r[0] == 'Line No': ix = {n: i for i, n in enumerate(r)}; continue
        if ix is None or not r[0].isdigit(): continue
        line = int(r[0])
        if cur not in maps: maps[cur] = func_map(cur)
        fn = "?"
        if maps[cur]:
            for s, name in maps[cur]:
                if s <= line: fn = name
                else: break
        def f(n):
            try: return float(r[ix[n]] or 0)
            except (ValueError, KeyError, IndexError): return 0.0
        c = agg[(cur, fn)]
        c['inst'] += f('Instructions Executed'); c['tinst'] += f('Thread Instructions Executed'); c['samples'] += f('# Samples')
        for n in ix:
            if n.startswith('stall_') and 'Not Issued' not in n: c[n] += f(n)
    tot = collections.Counter()
    for c in agg.values():
        for k, v in c.items(): tot[k] += v
    print("total warp inst %.4g, lanes/inst %.2f, samples %d" % (tot['inst'], tot['tinst'] / max(tot['inst'], 1), tot['samples']))
    print("%-44s %7s %6s %8s   dominant stalls" % ("function", "%inst", "lanes", "%samples"))
    for (fn, name), c in sorted(agg.items(), key=lambda kv: -kv[1]['samples'])[:topn]:
        st = sorted(((k[6:], v) for k, v in c.items() if k.startswith('stall_')), key=lambda t: -t[1])[:3]
        print("%-44s %6.2f%% %6.1f %7.2f%%   %s" % (fn[:-4] + ":" + name, 100 * c['inst'] / tot['inst'], c['tinst'] / max(c['inst'], 1), 100 * c['samples'] / max(tot['samples'], 1),
                                                    " ".join("%s %.0f%%" % (k, 100 * v / max(c['samples'], 1)) for k, v in st)))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 45)
