"""SASS / resource summary of libmt_b200.so per kernel: registers, stack, spills (LDL/STL), access widths, multimem.

    python tools/sass_summary.py [--all] > profiles/r2_sass_summary.txt

A spill inside the element loop shows up as LDL/STL between the first LDG and the last STG of a kernel ("hot" column:
counted only there); LDL/STL outside that window belong to prologue / slow paths (IEEE division helper, pow()).
"""
import re
import subprocess
import sys
from collections import defaultdict
from pathlib import Path

LIB = Path(__file__).resolve().parents[1] / "beforeholiday_b200" / "libmt_b200.so"


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def main():
    res = subprocess.run(["cuobjdump", "-res-usage", str(LIB)], capture_output=True, text=True).stdout
    usage = {}
    cur = None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            cur = m.group(1)
            continue
        if cur and "REG:" in line:
            usage[cur] = (int(re.search(r"REG:(\d+)", line).group(1)), int(re.search(r"STACK:(\d+)", line).group(1)))
            cur = None
    sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True).stdout
    stats = {}
    cur = None
    lines = []
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            if cur:
                stats[cur] = lines
            cur, lines = m.group(1), []
            continue
        if cur and re.search(r"/\*[0-9a-f]{4}\*/", line):
            lines.append(line)
    if cur:
        stats[cur] = lines
    dm = demangle(list(stats))
    rows = []
    for fn, ls in stats.items():
        ops = [re.sub(r"^\s*/\*[0-9a-f]+\*/\s*", "", l).strip() for l in ls]
        is_ldg = [bool(re.search(r"\bLDG", o)) for o in ops]
        is_stg = [bool(re.search(r"\bSTG", o)) for o in ops]
        first = is_ldg.index(True) if True in is_ldg else 0
        last = len(ops) - 1 - is_stg[::-1].index(True) if True in is_stg else len(ops)
        spill = [bool(re.search(r"\b(LDL|STL)", o)) for o in ops]
        # vector-path window: from the first 128/256-bit LDG to the last 128/256-bit STG (or last wide LDG for pure reductions)
        wide_ld = [i for i, o in enumerate(ops) if re.search(r"\bLDG\.[A-Z0-9.]*(128|256)", o) or "LDGMC" in o]
        wide_st = [i for i, o in enumerate(ops) if re.search(r"\bSTG\.[A-Z0-9.]*(128|256)", o)]
        hot = 0
        if wide_ld:
            lo, hi = wide_ld[0], (wide_st[-1] if wide_st else wide_ld[-1])
            hot = sum(spill[lo:hi + 1])
        name = dm.get(fn, fn)
        short = re.sub(r"^void mt::mt_table_kernel<mt::", "", name)
        short = re.sub(r"\(mt::DevTable.*$", "", short)
        reg, stack = usage.get(fn, (-1, -1))
        rows.append((short, reg, stack, sum(spill), hot,
                     sum(1 for o in ops if re.search(r"\bLDG\.[A-Z0-9.]*256", o)), sum(1 for o in ops if re.search(r"\bSTG\.[A-Z0-9.]*256", o)),
                     sum(1 for o in ops if re.search(r"\bLDG\.[A-Z0-9.]*128", o)), sum(1 for o in ops if re.search(r"\bSTG\.[A-Z0-9.]*128", o)),
                     sum(1 for o in ops if "LDGMC" in o), sum(1 for o in ops if re.search(r"\bCALL", o))))
    rows.sort()
    show_all = "--all" in sys.argv
    print("# libmt_b200.so (sm_100a): cuobjdump -res-usage / -sass per kernel.  'hot' = LDL+STL between the first wide LDG and the")
    print("# last wide STG (the vector element loop); LDL+STL outside it sit in the IEEE-division / pow() slow paths and peels.")
    print(f"# {'kernel':110s} regs stack LDL+STL hot LDG256 STG256 LDG128 STG128 LDGMC CALL")
    fam = defaultdict(list)
    for r in rows:
        fam[re.sub(r"<.*", "", r[0])].append(r)
        if show_all or r[3] > 0:
            print(f"{r[0][:112]:112s} {r[1]:4d} {r[2]:5d} {r[3]:7d} {r[4]:3d} {r[5]:6d} {r[6]:6d} {r[7]:6d} {r[8]:6d} {r[9]:5d} {r[10]:4d}")
    print("\n# per family: instantiations, regs min-max, total LDL+STL, total hot")
    for f, rs in sorted(fam.items()):
        print(f"{f:32s} {len(rs):3d}  {min(r[1] for r in rs)}-{max(r[1] for r in rs):3d}  {sum(r[3] for r in rs):5d} {sum(r[4] for r in rs):5d}")


if __name__ == "__main__":
    main()
