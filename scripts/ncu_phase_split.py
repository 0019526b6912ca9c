"""Per-source-line / per-phase split of a kernel's executed instructions from an ncu report captured with --import-source on.
   python scripts/ncu_phase_split.py report.ncu-rep kernel_regex source_file marker1 marker2 ...
Markers are substrings of source lines; a phase runs from its marker line to the next marker's line."""
import csv
import io
import subprocess
import sys


def load(rep, kernel):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kernel],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    per = {}; alu = {}; cur = None; fname = None
    hdr = None
    for r in rows:
        if len(r) >= 2 and r[0] == "File Path":
            fname = r[1]; continue
        if r and r[0] == "Line No":
            hdr = r; continue
        if hdr is None or len(r) < 8:
            continue
        if r[0] != "":
            try:
                cur = (fname, int(r[0]))
            except ValueError:
                cur = None
            continue
        ie = hdr.index("Instructions Executed")
        try:
            n = int(r[ie])
        except ValueError:
            continue
        if cur is None:
            continue
        per[cur] = per.get(cur, 0) + n
        op = r[3].strip().split()[0] if r[3].strip() else ""
        if op.startswith("@"):
            op = r[3].strip().split()[1]
        base = op.split(".")[0]
        if base in ("LOP3", "PRMT", "SHF", "VIMNMX", "VIMNMX3", "IADD3", "ISETP", "LEA", "VIADD", "IMNMX", "SEL", "POPC", "FLO", "BREV", "PLOP3", "VABSDIFF", "VABSDIFF4", "SGXT", "BMSK", "IABS"):
            alu[cur] = alu.get(cur, 0) + n
    return per, alu


if __name__ == "__main__":
    rep, kernel, srcfile = sys.argv[1:4]
    marks = sys.argv[4:]
    per, alu = load(rep, kernel)
    tot = sum(per.values())
    src = open(srcfile).read().split("\n")
    files = sorted({f for f, _ in per})
    print("total warp instructions: %d; source files: %s" % (tot, files))
    main = [f for f in files if f.endswith(srcfile.split("/")[-1])]
    mainf = main[0] if main else files[0]
    lines = {ln: n for (f, ln), n in per.items() if f == mainf}
    alines = {ln: n for (f, ln), n in alu.items() if f == mainf}
    other = sum(n for (f, ln), n in per.items() if f != mainf)
    pos = []
    for m in marks:
        for i, l in enumerate(src):
            if m in l:
                pos.append((m, i + 1)); break
    pos.append(("<end>", len(src) + 1))
    print("other files (inlined headers): %d (%.1f%%)" % (other, 100.0 * other / tot))
    first = pos[0][1]
    s = sum(n for ln, n in lines.items() if ln < first); a = sum(n for ln, n in alines.items() if ln < first)
    print("%-50s %12d %5.1f%%   alu %5.1f%%" % ("<before first marker: helpers>", s, 100.0 * s / tot, 100.0 * a / tot))
    for (m, a0), (_, b0) in zip(pos, pos[1:]):
        s = sum(n for ln, n in lines.items() if a0 <= ln < b0); a = sum(n for ln, n in alines.items() if a0 <= ln < b0)
        print("%-50s %12d %5.1f%%   alu %5.1f%%" % (m[:50], s, 100.0 * s / tot, 100.0 * a / tot))
    print("-- top lines")
    for ln, n in sorted(lines.items(), key=lambda x: -x[1])[:30]:
        print("%5d %11d %5.1f%%  %s" % (ln, n, 100.0 * n / tot, src[ln - 1].strip()[:120]))
