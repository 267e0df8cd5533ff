"""Per-source-line table of ONE kernel of an .ncu-rep captured with --import-source on (-lineinfo build):
share of warp instructions, lanes per instruction, share of stall samples; per file first, then the top lines.
usage: ncu_lines_kernel.py report.ncu-rep <kernel regex> [top] [sort: inst|samples]"""
import collections, csv, io, subprocess, sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
by = sys.argv[4] if len(sys.argv) > 4 else "inst"
out = subprocess.run(["ncu", "-i", rep, "-k", "regex:" + kern, "--page", "source", "--csv", "--print-source",
                      "cuda,sass"], capture_output=True, text=True).stdout
cur, res = None, []
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    try:
        ln, samples, inst, tinst = int(r[0]), int(r[6]), int(r[7]), int(r[8])
    except (ValueError, IndexError):
        continue
    res.append((cur, ln, r[1].strip()[:100], samples, inst, tinst))
ti, tt, ts = (sum(x[k] for x in res) for k in (4, 5, 3))
print(f"{kern}: {ti} warp instructions, {tt / max(ti, 1):.2f} lanes per instruction, {ts} stall samples")
files = collections.defaultdict(lambda: [0, 0, 0])
for f, _, _, s, i, t in res:
    a = files[f]
    a[0] += s
    a[1] += i
    a[2] += t
for f, a in sorted(files.items(), key=lambda x: -x[1][1])[:14]:
    print(f"  {f:28s} inst {100 * a[1] / ti:5.1f}%  lanes {a[2] / max(a[1], 1):5.1f}  samples {100 * a[0] / max(ts, 1):5.1f}%")
res.sort(key=lambda x: -(x[3] if by == "samples" else x[4]))
for f, ln, src, s, i, t in res[:top]:
    print(f"{f:16s}{ln:5d} inst {100 * i / ti:5.1f}% lanes {t / max(i, 1):5.1f} smp {100 * s / max(ts, 1):5.1f}% | {src}")
