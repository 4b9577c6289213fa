"""Per-kernel SASS instruction counts of the built objects (pinot_b200/csrc/build/*.o) -> profiles/r02/sass_summary.txt.

    python scripts/sass_summary.py > profiles/r02/sass_summary.txt

Runs on the CPU box (cuobjdump needs no GPU).  The opcodes that prove the tensor-core / TMA path are the ones
/opt/skills/guides/B200_PROFILING.md names: UTCHMMA (tcgen05.mma), UTMALDG (TMA load), LDTM (tcgen05.ld)."""
import glob, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COLS = [("UTCHMMA", r"UTC[HQ]MMA"), ("UTMALDG", r"UTMALDG"), ("UTMAPF", r"UTMAPF"), ("LDTM", r"LDTM"), ("UTCBAR", r"UTCBAR"),
        ("SYNCS", r"SYNCS"), ("FFMA", r"FFMA"), ("MUFU", r"MUFU"), ("LDL", r"LDL"), ("STL", r"STL")]


def demangle(names):
    out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.split("\n")
    clean = []
    for o in out[:len(names)]:
        o = re.sub(r"^void ", "", o).replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
        o = re.sub(r"\((int|bool|unsigned int|long)\)", "", o)
        clean.append(re.sub(r"\(.*$", "", o).replace("pb::", ""))
    return clean


print("# cuobjdump -sass of pinot_b200/csrc/build/*.o (sm_100a), round-2 code: instruction counts per kernel (scripts/sass_summary.py)")
print("# UTCHMMA (incl. UTCQMMA) = tcgen05.mma, UTMALDG = TMA load, UTMAPF = TMA prefetch into L2, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit,")
print("# SYNCS = mbarrier ops; FFMA / MUFU = CUDA-core arithmetic; LDL / STL = local memory")
print("%-92s" % "kernel" + "".join("%8s" % c for c, _ in COLS) + "%8s" % "total")
for obj in sorted(glob.glob(os.path.join(ROOT, "pinot_b200", "csrc", "build", "*.o"))):
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    kernels, cur = [], None
    for line in sass.split("\n"):
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = {"name": m.group(1), "n": 0, **{c: 0 for c, _ in COLS}}
            kernels.append(cur)
            continue
        if cur is None or not re.match(r"\s+/\*[0-9a-f]{4,}\*/\s", line):
            continue
        cur["n"] += 1
        for c, pat in COLS:
            if re.search(r"\b" + pat, line):
                cur[c] += 1
    if not kernels:
        continue
    print("## " + os.path.basename(obj))
    for k, name in zip(kernels, demangle([k["name"] for k in kernels])):
        print("%-92s" % name[:91] + "".join("%8d" % k[c] for c, _ in COLS) + "%8d" % k["n"])
