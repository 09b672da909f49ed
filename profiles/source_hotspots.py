"""Per-source-line instruction and stall totals from an `ncu --page source --print-source cuda,sass --csv` export.

    ncu -i capture.ncu-rep --page source --print-source cuda,sass --csv > capture_sc.csv
    python profiles/source_hotspots.py capture_sc.csv [top]

Rows with a line number are CUDA source lines (metrics summed over their SASS); prints, for list_metrics.cu,
the executed warp instructions, the share of the kernel, stall samples and shared-memory wavefronts per line.
"""
import csv
import sys


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    cur_file, hdr, lines = None, None, []
    for r in csv.reader(open(path, errors="replace")):
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1]
        elif r[0] == "Line No":
            hdr = r
        elif hdr and r[0].isdigit() and len(r) >= len(hdr):
            # source text with quotes and commas in it splits into extra fields: align the metrics from the right
            extra = len(r) - len(hdr)
            d = dict(zip(hdr[2:], r[2 + extra:]))
            lines.append((cur_file, int(r[0]), ",".join(r[1:2 + extra]), d))
    num = lambda d, k: float(d.get(k, "0").replace(",", "") or 0) if d.get(k, "0") not in ("-", "") else 0.0
    tot = sum(num(d, "Instructions Executed") for _, _, _, d in lines)
    samp = sum(num(d, "# Samples") for _, _, _, d in lines)
    wav = sum(num(d, "L1 Wavefronts Shared") for _, _, _, d in lines)
    print("total warp instructions %d, samples %d, shared wavefronts %d" % (tot, samp, wav))
    lines.sort(key=lambda x: -num(x[3], "Instructions Executed"))
    print("%-22s %10s %6s %7s %7s %9s %8s  source" % ("file:line", "inst", "%", "samp%", "short_sb", "wavefr", "excess"))
    for f, n, src, d in lines[:top]:
        print("%-22s %10d %6.2f %7.2f %7.2f %9d %8d  %s" % (
            f.split("/")[-1] + ":" + str(n), num(d, "Instructions Executed"), 100 * num(d, "Instructions Executed") / tot,
            100 * num(d, "# Samples") / max(samp, 1), 100 * num(d, "stall_short_sb") / max(samp, 1),
            num(d, "L1 Wavefronts Shared"), num(d, "L1 Wavefronts Shared Excessive"), src.strip()[:90]))


if __name__ == "__main__":
    main()
