"""Where are the spills?  Prints, for one kernel of an object file, the SASS line ranges of local
loads/stores next to role markers (USETMAXREG, first/last MUFU.EX2, UTCHMMA, UTMALDG, backward branches).
  python tools/sass_spills.py <obj> <mangled-kernel-substring>"""
import re
import subprocess
import sys

obj, pat = sys.argv[1], sys.argv[2]
names = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
cur, lines = None, []
for ln in names.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = m.group(1)
        continue
    if cur and pat in cur and re.search(r"/\*[0-9a-f]{4,}\*/", ln):
        lines.append(ln)
ins = []
for ln in lines:
    m = re.search(r"/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
print(f"{len(ins)} instructions")
marks = []
for i, (pc, t) in enumerate(ins):
    op = t.split()[1] if t.startswith("@") else t.split()[0]
    kind = None
    if op.startswith(("STL", "LDL")): kind = op.split(".")[0]
    elif op.startswith("USETMAXREG"): kind = t
    elif op.startswith("MUFU.EX2"): kind = "EX2"
    elif op.startswith("UTCHMMA"): kind = "MMA"
    elif op.startswith("UTMALDG"): kind = "TMA"
    elif op.startswith("BRA"):
        m = re.search(r"0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) < pc: kind = f"LOOP-BACK->{int(m.group(1), 16) // 16}"
    if kind: marks.append((i, kind))
# run-length compress
out, j = [], 0
while j < len(marks):
    k = j
    while k + 1 < len(marks) and marks[k + 1][1] == marks[j][1] and not marks[j][1].startswith(("LOOP", "USET")): k += 1
    out.append(f"{marks[j][0]}" + (f"-{marks[k][0]}" if k > j else "") + f":{marks[j][1]}" + (f"x{k - j + 1}" if k > j else ""))
    j = k + 1
print("  ".join(out))
