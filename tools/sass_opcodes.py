"""SASS opcode counts per kernel of libmplb200.so -> profiles/rN_sass_opcodes.txt (the mnemonics that prove the hardware paths:
UTCHMMA = tcgen05.mma kind::f16 / tf32, UTCQMMA = kind::f8f6f4, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UTMALDG = TMA tile copy, UBLKCP = 1-D TMA bulk copy, SYNCS =
mbarrier, POPC / LOP3 / VIMNMX = the Hamming integer path, DFMA / DMUL / DADD = fp64, LDL / STL = local memory).
    python tools/sass_opcodes.py profiles/r2_sass_opcodes.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "matchinglib_poselib_b200", "lib", "libmplb200.so")
OPS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "UTMALDG", "UBLKCP", "SYNCS", "POPC", "LOP3", "VIMNMX", "DFMA", "DMUL", "DADD", "MUFU", "LDL", "STL",
       "ATOM", "ATOMS", "ATOMG", "RED", "SHFL", "VOTE", "UCGABAR_ARV"]


def main(out_path):
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    funcs, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1); funcs[cur] = collections.Counter(); continue
        if cur:
            m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
            if m:
                funcs[cur][m.group(1).split(".")[0]] += 1

    def demangle(n):
        r = subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
        return re.sub(r"\(.*", "", re.sub(r"\(anonymous namespace\)::", "", r))
    with open(out_path, "w") as f:
        f.write("# " + __doc__.strip().replace("\n", "\n# ") + "\n")
        f.write("%-44s %6s " % ("kernel", "instrs") + " ".join("%7s" % o for o in OPS) + "\n")
        for n, c in funcs.items():
            f.write("%-44s %6d " % (demangle(n)[:44], sum(c.values())) + " ".join("%7d" % c.get(o, 0) for o in OPS) + "\n")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r2_sass_opcodes.txt"))
