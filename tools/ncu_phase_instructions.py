"""Executed warp instructions of a kernel attributed to the CALLING code: the inlined vector arithmetic of sim_math.cuh (a third
of all executed instructions) is charged to the nearest preceding instruction, in address order, that comes from another file
(instructions of one inlined call are contiguous), and the result is summed per 20-line window of each file.
   python tools/ncu_phase_instructions.py x.ncu-rep [topN]"""
import collections
import csv
import subprocess
import sys

LEAF = ("sim_math.cuh", "sm_30_intrinsics.hpp", "sm_32_intrinsics.hpp", "device_atomic_functions.hpp", "math_functions.hpp", "device_functions.hpp")


def main(rep, topn=60):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    cur = None; ix = None; line = None
    ins = []      # (addr, file, line, executed, thread executed)
    src = {}
    for r in csv.reader(out.splitlines()):
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]; continue
        if r[0] == "Line No":
            ix = {n: i for i, n in enumerate(r)}
            ie, it = ix["Instructions Executed"], ix["Thread Instructions Executed"]
            continue
        if ix is None or r[0] == "Function Name":
            continue
        if r[0] == "":
            try:
                ins.append((int(r[2], 16), cur, line, int(float(r[ie] or 0)), int(float(r[it] or 0))))
            except ValueError:
                pass
            continue
        if r[0].isdigit():
            line = int(r[0]); src[(cur, line)] = r[1].strip()[:90]
    ins.sort()
    tot = sum(x[3] for x in ins)
    ctx = ("?", 0)
    B = collections.defaultdict(lambda: [0, 0, collections.Counter()])
    for a, f, l, e, t in ins:
        if f not in LEAF:
            ctx = (f, l)
        b = B[(ctx[0], ctx[1] // 20 * 20)]
        b[0] += e; b[1] += t; b[2][ctx] += e
    print("executed warp instructions %d; share / lanes per instruction / window / its heaviest line" % tot)
    for (f, l), (e, t, c) in sorted(B.items(), key=lambda kv: -kv[1][0])[:topn]:
        best = c.most_common(1)[0][0]
        print("  %5.2f %%  %5.1f  %-16s %4d-%-4d  %s" % (100.0 * e / tot, t / max(1, e), f, l, l + 19, src.get(best, "")))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 60)
