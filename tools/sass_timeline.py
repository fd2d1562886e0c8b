"""Issue-order view of one kernel's SASS with the scheduling control fields decoded (development tool).
  nvdisasm -hex <cubin> > all.sass ; python tools/sass_timeline.py all.sass <mangled-name-substring> [first_line last_line]
Per instruction: index, stall count, yield, write/read scoreboard, wait mask, text.  With a line range it also
prints the sum of the stall fields (a lower bound of a lone warp's issue time over straight-line code) and the
per-opcode histogram.  Control word layout (bits of the 128-bit instruction): stall [105,109), yield 109,
wbar [110,113), rbar [113,116), wait mask [116,122)."""
import collections
import re
import sys

path, pat = sys.argv[1], sys.argv[2]
lo = int(sys.argv[3]) if len(sys.argv) > 3 else None
hi = int(sys.argv[4]) if len(sys.argv) > 4 else None
lines = open(path).read().split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith(".text.") and pat in l and l.rstrip().endswith(":"))
ins = []
cur = None
labels = {}
for ln in lines[start + 1:]:
    if ln.startswith("//-----"):
        break
    ml = re.match(r"(\.L_x_\d+):", ln)
    if ml:
        labels[len(ins)] = ml.group(1)
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s+/\* (0x[0-9a-f]{16}) \*/", ln)
    if m:
        cur = [int(m.group(1), 16), m.group(2).strip(), int(m.group(3), 16), None]
        continue
    m = re.match(r"\s+/\* (0x[0-9a-f]{16}) \*/", ln)
    if m and cur:
        cur[3] = int(m.group(1), 16)
        ins.append(cur)
        cur = None
rows = []
for i, (pc, text, lo64, hi64) in enumerate(ins):
    c = hi64 >> 41
    stall, yld, wbar, rbar, wait = c & 0xf, (c >> 4) & 1, (c >> 5) & 7, (c >> 8) & 7, (c >> 11) & 0x3f
    rows.append((i, pc, stall, yld, wbar, rbar, wait, text))
sel = rows if lo is None else [r for r in rows if lo <= r[0] <= (hi if hi is not None else lo + 400)]
tot, hist = 0, collections.Counter()
for i, pc, stall, yld, wbar, rbar, wait, text in sel:
    w = "".join(str(b) for b in range(6) if wait >> b & 1)
    if i in labels:
        print(f"{labels[i]}:")
    print(f"{i:5d} {pc:6x} s{stall:2d} {'Y' if yld else ' '} w{wbar if wbar < 7 else '-'} r{rbar if rbar < 7 else '-'} wait[{w:3s}] {text}")
    tot += stall
    op = text.split()[1] if text.startswith("@") else text.split()[0]
    hist[op.split(".")[0]] += 1
if lo is not None:
    print(f"sum of stall fields: {tot} over {len(sel)} instructions")
    print(dict(hist.most_common()))
