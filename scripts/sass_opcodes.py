"""Per-kernel counts of the SASS mnemonics that prove (or disprove) a Blackwell-native kernel:
UTCHMMA (tcgen05.mma), UTMALDG / UTMASTG (TMA load / store), LDTM / STTM (tcgen05.ld / st), HMMA (legacy mma.sync),
MUFU.EX2, LDGSTS (cp.async), RED/ATOM.  Reads the in-tree .so with cuobjdump; writes profiles/sass_opcodes.txt.
Usage: python scripts/sass_opcodes.py [out]"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "transformer-cnn-hybrid-network-for-video-processing_b200", "libhybridnet_sm100.so")
OPS = ["UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "STTM", "HMMA", "MUFU.EX2", "LDGSTS", "MOVM", "RED", "ATOMG", "ATOMS"]


def main(out):
    txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    counts, name = collections.OrderedDict(), None
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = name.replace("(anonymous namespace)::", "").replace("void ", "")
            name = re.sub(r"\(.*", "", name)
            counts[name] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and name:
            op = m.group(1)
            for o in OPS:
                if op == o or op.startswith(o + "."):
                    counts[name][o] += 1
            counts[name]["_all"] += 1
    with open(out, "w") as f:
        f.write("SASS opcode summary of libhybridnet_sm100.so (cuobjdump -sass; static instruction counts per kernel)\n")
        f.write(f"{'kernel':72s} {'instr':>6s} " + " ".join(f"{o:>8s}" for o in OPS) + "\n")
        for k, c in counts.items():
            if c["_all"] == 0:
                continue
            f.write(f"{k[:72]:72s} {c['_all']:6d} " + " ".join(f"{c[o]:8d}" for o in OPS) + "\n")
    print("wrote", out)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "sass_opcodes.txt"))
