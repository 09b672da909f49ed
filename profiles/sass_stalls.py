"""Stall mix and the most-stalled SASS instructions of one kernel, from
`ncu -i capture.ncu-rep --page source --print-source sass --csv > capture_sass.csv`.

    python profiles/sass_stalls.py capture_sass.csv [top] [context]
"""
import csv
import sys


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 12
    ctx = int(sys.argv[3]) if len(sys.argv) > 3 else 6
    rows = list(csv.reader(open(path, errors="replace")))
    hdr, body = rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    f = lambda r, i: float(r[i]) if i < len(r) and r[i].replace(".", "").isdigit() else 0.0
    samples = sum(f(r, col["# Samples"]) for r in body)
    inst = sum(f(r, col["Instructions Executed"]) for r in body)
    print("warp instructions %d, samples %d" % (inst, samples))
    mix = {hdr[i]: sum(f(r, i) for r in body) for i in stall}
    tot = sum(mix.values())
    print("stall mix: " + ", ".join("%s %.1f%%" % (k[6:], 100 * v / tot) for k, v in sorted(mix.items(), key=lambda x: -x[1])[:9]))
    order = sorted(range(len(body)), key=lambda i: -f(body[i], col["# Samples"]))[:top]
    for t in order:
        r = body[t]
        why = sorted(((f(r, i), hdr[i][6:]) for i in stall), reverse=True)[:2]
        print("---- %.2f%% of samples: %s" % (100 * f(r, col["# Samples"]) / samples, ", ".join("%s %d" % (n, v) for v, n in why)))
        for q in body[max(0, t - ctx):t + 2]:
            print("     %s%s" % ("> " if q is r else "  ", q[col["Source"]].strip()[:100]))


if __name__ == "__main__":
    main()
