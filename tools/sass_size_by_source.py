"""Static code size of env_step_kernel by source file / line (nvdisasm -g on the cubin of build/envset.o): how many SASS
instructions each source line contributes, split into the kernel's main body and its out-of-line device functions. The
kernel is instruction-fetch bound (profiles/r2b_*), so code size along the per-tick path is what is optimised.
    python tools/sass_size_by_source.py [envset.o] [topN]"""
import collections
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    obj = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "reinforcement-learning_b200", "build", "envset.o")
    topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
    cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    out = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
    in_k = False
    cur = ("?", 0)
    by_line = collections.Counter(); by_file = collections.Counter()
    addr_file = []   # (addr, file)
    n = 0
    for line in out.splitlines():
        if line.startswith("//---") and ".text." in line:
            in_k = "env_step_kernel" in line
            continue
        if not in_k:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', line)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/", line)
        if m:
            n += 1
            by_line[cur] += 1; by_file[cur[0]] += 1
            addr_file.append((int(m.group(1), 16), cur[0]))
    print("env_step_kernel: %d SASS instructions = %.0f KB" % (n, n * 16 / 1024))
    print("\nby file:")
    for f, c in by_file.most_common():
        print("  %-20s %6d  %5.1f %%" % (f, c, 100.0 * c / n))
    print("\ntop lines:")
    srcs = {}
    for (f, l), c in by_line.most_common(topn):
        p = os.path.join(ROOT, "reinforcement-learning_b200", "csrc", f)
        if f not in srcs and os.path.exists(p):
            srcs[f] = open(p).read().splitlines()
        text = srcs.get(f, [""] * (l + 1))[l - 1].strip()[:100] if f in srcs and l - 1 < len(srcs[f]) else ""
        print("  %5d  %s:%d  %s" % (c, f, l, text))
    # address map: 4 KB buckets -> dominant file
    print("\naddress map (16 KB buckets: dominant files):")
    buckets = collections.defaultdict(collections.Counter)
    for a, f in addr_file:
        buckets[a // 16384][f] += 1
    for b in sorted(buckets):
        top = ", ".join("%s %d" % (f, c) for f, c in buckets[b].most_common(3))
        print("  %6x  %s" % (b * 16384, top))


if __name__ == "__main__":
    main()
