"""Tensor-core / TMA / TMEM / mbarrier SASS mnemonics per kernel: `cuobjdump -sass` of gsc_b200/build/*.o.
Usage: python tools/sass_summary.py > profiles/r2_sass_summary.txt   (no GPU needed)"""
import collections, glob, os, re, subprocess
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pat = re.compile(r"\b(UTCHMMA|UTCBAR|LDTM|STTM|UTCATOMSWS|UTMALDG|UTMASTG|ELECT|FENCE\.VIEW\.ASYNC\.[ST]|LDGSTS|"
                 r"SYNCS\.[A-Z0-9.]+|HMMA\.1688\.F32\.TF32|MATCH\.ANY)\b")
print("# Tensor-core / TMA / TMEM / mbarrier SASS mnemonics per kernel (cuobjdump -sass of gsc_b200/build/*.o, sm_100a)")
print("# UTCHMMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM / STTM = tcgen05.ld / st, UTCATOMSWS = tcgen05.alloc/dealloc,")
print("# UTMALDG / UTMASTG = cp.async.bulk.tensor load / store, SYNCS.* = mbarrier ops, HMMA.1688.F32.TF32 = mma.sync m16n8k8 tf32")
for obj in sorted(glob.glob(os.path.join(root, "gsc_b200", "build", "*.o"))):
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    name, counts = None, None
    def flush():
        if name and counts:
            dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
            dem = re.sub(r"\(.*", "", dem)
            print(f"{dem}: " + ", ".join(f"{k} x{v}" for k, v in sorted(counts.items())))
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            flush()
            name, counts = m.group(1), collections.Counter()
            continue
        if counts is not None:
            for tok in pat.findall(line):
                counts[tok] += 1
    flush()
