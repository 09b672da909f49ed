"""Writes profiles/sass_markers.txt: how often the Blackwell-native SASS mnemonics occur in each kernel of
libflr_b200.so (cuobjdump -sass; run here, no GPU needed).   python profiles/sass_markers.py"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "fairness-llm-replication_b200", "libflr_b200.so")
MARKS = ["UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "SYNCS", "ATOMS", "MATCH", "REDUX", "HMMA"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    cur, counts = None, collections.OrderedDict()
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
        elif cur:
            for k in MARKS:
                if re.search(r"\b" + k + r"\b", line):
                    counts[cur][k] += 1
    names = subprocess.run(["cu++filt"] + list(counts), capture_output=True, text=True).stdout.splitlines()
    total = collections.Counter()
    with open(os.path.join(ROOT, "profiles", "sass_markers.txt"), "w") as f:
        f.write("# cuobjdump -sass libflr_b200.so (sm_100a cubin only): occurrences of Blackwell-native mnemonics per kernel.\n"
                "# UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG = TMA tensor load/store,\n"
                "# UBLKCP = cp.async.bulk (1-D TMA), UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, ATOMS = shared-memory\n"
                "# atomics; HMMA (the legacy tensor path) must stay 0.  Kernels with none of them are omitted.\n")
        for (mangled, c), name in zip(counts.items(), names):
            name = re.sub(r"\(anonymous namespace\)::|<unnamed>::|void ", "", name)
            depth, cut = 0, len(name)
            for i, ch in enumerate(name):          # drop the parameter list, keep template arguments
                if ch == "<":
                    depth += 1
                elif ch == ">":
                    depth -= 1
                elif ch == "(" and depth == 0:
                    cut = i
                    break
            name = name[:cut].replace("(int)", "").replace("(bool)", "")
            if sum(c.values()):
                f.write("%-56s %s\n" % (name, " ".join("%s=%d" % (m, c[m]) for m in MARKS if c[m])))
            total.update(c)
        f.write("TOTAL %s\n" % " ".join("%s=%d" % (m, total[m]) for m in MARKS))


if __name__ == "__main__":
    main()
