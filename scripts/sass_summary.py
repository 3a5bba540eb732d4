"""Per-kernel SASS evidence for the built library: counts of the tensor-core / TMEM / TMA mnemonics that prove which pipe a kernel
uses (UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG = cp.async.bulk.tensor load / store, UTCBAR =
tcgen05.commit, HMMA = mma.sync, UBLKCP = cp.async.bulk), plus registers and spill bytes from the ptxas logs.
    python scripts/sass_summary.py > profiles/sass_summary.txt          (needs cuobjdump; no GPU)"""
import collections
import hashlib
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "dualhyp_b200", "libdualhyp.so")
MNEMONICS = ["UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "HMMA", "UBLKCP", "LDGSTS", "LDSM", "REDG", "STL", "LDL"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return [re.sub(r"\(.*", "", n).replace("void ", "").replace("dh::", "") for n in out]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            op = m.group(1)
            counts[cur]["insts"] += 1
            for k in MNEMONICS:
                if op.startswith(k):
                    counts[cur][k] += 1
    regs = {}
    logdir = os.path.join(ROOT, "dualhyp_b200", "csrc", "build")
    for f in sorted(os.listdir(logdir)) if os.path.isdir(logdir) else []:
        if not f.endswith(".ptxas.log"):
            continue
        txt = open(os.path.join(logdir, f)).read()
        for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'.*?(\d+) bytes spill stores.*?Used (\d+) registers", txt, re.S):
            regs[m.group(1)] = (int(m.group(3)), int(m.group(2)))
    names = demangle(list(counts))
    sha = hashlib.sha256(open(LIB, "rb").read()).hexdigest()[:16]
    print(f"libdualhyp.so sha256/16 {sha}; cuobjdump -sass, sm_100a; columns: SASS instructions, registers, spill-store bytes, then mnemonic counts")
    hdr = f"{'kernel':64s} {'insts':>6s} {'regs':>4s} {'spill':>5s} " + " ".join(f"{k:>7s}" for k in MNEMONICS)
    print(hdr)
    tot = collections.Counter()
    for (raw, c), name in sorted(zip(counts.items(), names), key=lambda x: x[1]):
        r = regs.get(raw, (0, 0))
        print(f"{name[:64]:64s} {c['insts']:6d} {r[0]:4d} {r[1]:5d} " + " ".join(f"{c[k]:7d}" for k in MNEMONICS))
        tot.update(c)
    print(f"{'TOTAL':64s} {tot['insts']:6d} {'':4s} {'':5s} " + " ".join(f"{tot[k]:7d}" for k in MNEMONICS))


if __name__ == "__main__":
    sys.exit(main())
