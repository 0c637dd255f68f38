"""SASS mnemonic counts per kernel of the in-tree library (no GPU needed): the proof that the hot kernels issue
tcgen05 / TMA / mbarrier instructions.

    python benchmarks/sass_summary.py > profiles/sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pytorchdistributed_b200", "lib", "libptdb200.so")
KEEP = re.compile(r"^(UTC|UTMA|LDTM|STTM|SYNCS|LDGSTS|LDGMC|RED|REDG|UBLKCP|UBLKPF|UCGABAR|ATOM|CCTL|ELECT|ACQBULK|PREEXIT)|\.SYS")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    names = sass
    per = collections.OrderedDict()
    cur = None
    for line in names.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            per[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", line)
        if m and cur is not None and KEEP.search(m.group(1)):
            per[cur][m.group(1)] += 1
    demangled = subprocess.run(["c++filt"] + list(per.keys()), capture_output=True, text=True).stdout.splitlines()
    print("SASS mnemonic counts per kernel (cuobjdump -sass libptdb200.so; sm_100a).")
    print("UTCHMMA = tcgen05.mma (.2CTA = cta_group::2), LDTM = tcgen05.ld, UTMALDG(.IM2COL) = cp.async.bulk.tensor load, "
          "UTMASTG = TMA store,\nUTCBAR = tcgen05.commit, SYNCS = mbarrier, LDGSTS = cp.async, RED/REDG = red.global, "
          "UCGABAR = cluster barrier, *.SYS loads/stores = peer (NVLink) accesses in the collectives,\nLDGMC = multimem.ld_reduce (the NVSwitch reduction; multimem.st / multimem.red show up as STG / REDG on the multicast address),\n"
          "ACQBULK = griddepcontrol.wait, PREEXIT = griddepcontrol.launch_dependents (programmatic dependent launch).\n")
    rows = []
    for (mangled, cnt), name in zip(per.items(), demangled):
        if not cnt:
            continue
        short = name.replace("(anonymous namespace)::", "").replace("void ", "")
        cut = short.find(">(")
        short = short[:cut + 1] if cut >= 0 else short.split("(")[0]
        short = re.sub(r"\((?:int|bool|unsigned int)\)", "", short)
        rows.append((short, ", ".join(f"{k} x{v}" for k, v in sorted(cnt.items()))))
    for short, body in sorted(rows):
        print(short)
        print("    " + body)


if __name__ == "__main__":
    sys.exit(main())
