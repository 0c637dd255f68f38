"""SASS LISTINGS (not just mnemonic counts) of the hot loops, cut out of `cuobjdump -sass` of the in-tree library:

    python benchmarks/sass_listing.py > profiles/sass_listing.txt

For each kernel named below: the instructions around its tensor-core / TMA / multimem instructions (a window of
context lines before and after every match, merged), with addresses - enough to read the main loop.
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pytorchdistributed_b200", "lib", "libptdb200.so")
# (demangled-name pattern, instructions that mark the interesting region, context lines)
WANT = [
    (r"igemm_f_kernel<256, 2, false, 0, false, false>", r"UTCHMMA|UTMALDG|UTCBAR|LDTM|UTMASTG|ACQBULK|PREEXIT", 6),
    (r"igemm_f_kernel<256, 1, true, 0, true, false>", r"UTCHMMA|UTMALDG|UTCBAR", 6),          # CTA pair (cta_group::2), dgrad form
    (r"igemm_w_kernel<4, 2, true>", r"UTCHMMA|UTMALDG|UTCBAR|REDG|RED\.", 5),
    (r"allreduce_mc_kernel<false>", r"LDGMC|STG\.E\.128|REDG|MEMBAR|LDG\.E\.STRONG", 4),
    (r"allreduce_wire_bf16_kernel<8, true>", r"LDGMC|STG\.E\.128\.STRONG|REDG", 3),
    (r"broadcast_mc_kernel", r"STG\.E\.128\.STRONG|REDG", 3),
]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    funcs, cur, name = {}, None, None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            cur = funcs.setdefault(name, [])
            continue
        if cur is not None and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
            cur.append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", line.rstrip()))
    mangled = list(funcs.keys())
    demangled = subprocess.run(["c++filt"] + mangled, capture_output=True, text=True).stdout.splitlines()
    print("SASS listings of the hot loops (cuobjdump -sass libptdb200.so, sm_100a; encodings stripped).")
    print("UTCHMMA = tcgen05.mma, UTMALDG = TMA load, UTCBAR = tcgen05.commit, LDTM = tcgen05.ld, UTMASTG = TMA store,")
    print("LDGMC = multimem.ld_reduce, STG/REDG on the multicast address = multimem.st / multimem.red.\n")
    for pat, marks, ctx in WANT:
        hits = [(m, d) for m, d in zip(mangled, demangled) if pat in d.replace("(anonymous namespace)::", "")]
        if not hits:
            print(f"== {pat}: not found in this build\n")
            continue
        m, d = hits[0]
        lines = funcs[m]
        keep = set()
        for i, ln in enumerate(lines):
            if re.search(marks, ln):
                keep.update(range(max(0, i - ctx), min(len(lines), i + ctx + 1)))
        print(f"== {d.replace('(anonymous namespace)::', '').split('(')[0]}   ({len(lines)} instructions, {len(keep)} shown)")
        last = -2
        for i in sorted(keep):
            if i != last + 1:
                print("        ...")
            print(lines[i])
            last = i
        print()


if __name__ == "__main__":
    sys.exit(main())
