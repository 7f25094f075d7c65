"""Per-kernel count of the Blackwell-specific SASS mnemonics in libb200sl.so (cuobjdump -sass), written to
profiles/<name>: UTCHMMA (tcgen05.mma; .2CTA = cta_group::2), LDTM / STTM (tcgen05.ld / st), UTMALDG / UTMASTG / UTMAREDG
(TMA load / store / reduce-add), UTCBAR (tcgen05.commit), SYNCS (mbarrier), plus HMMA (legacy mma.sync: expected 0)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "learning-jax-sharding_b200", "libb200sl.so")
OPS = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTCBAR", "SYNCS", "HMMA", "FFMA", "MUFU.EX2"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts = collections.OrderedDict()
    name = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
            counts.setdefault(name, collections.Counter())
            continue
        if name is None:
            continue
        m = re.search(r"^\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        for key in OPS:
            if op == key or op.startswith(key + "."):
                counts[name][key] += 1
    lines = ["kernel".ljust(64) + "".join(k.rjust(14) for k in OPS)]
    total = collections.Counter()
    for kname, c in counts.items():
        lines.append(kname[:63].ljust(64) + "".join(str(c[k]).rjust(14) for k in OPS))
        total.update(c)
    lines.append("TOTAL".ljust(64) + "".join(str(total[k]).rjust(14) for k in OPS))
    text = "\n".join(lines) + "\n"
    dst = os.path.join(ROOT, "profiles", sys.argv[1] if len(sys.argv) > 1 else "sass_mnemonics.txt")
    with open(dst, "w") as f:
        f.write("cuobjdump -sass learning-jax-sharding_b200/libb200sl.so (sm_100a), instruction counts per kernel (UTCHMMA counts every tcgen05.mma, UTCHMMA.2CTA the cta_group::2 ones among them)\n" + text)
    print(text)


if __name__ == "__main__":
    main()
