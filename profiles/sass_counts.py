"""SASS mnemonic counts per kernel family of speechbrain_b200/csrc/libsbfe.so (cuobjdump -sass):
the evidence for packed f32x2 math, tensor-core instructions (mma.sync and tcgen05), cp.async and
programmatic dependent launch.   python profiles/sass_counts.py > profiles/r2_sass_counts.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "speechbrain_b200", "csrc", "libsbfe.so")
COLS = [("instructions", r"."),
        ("FFMA2/FADD2/FMUL2 (packed f32x2)", r"^(FFMA2|FADD2|FMUL2)"),
        ("HMMA.1688.F32.TF32 (mma.sync)", r"^HMMA"),
        ("UTCHMMA (tcgen05.mma)", r"^UTC[A-Z]*MMA"),
        ("STTM / LDTM (tcgen05.st / ld)", r"^(STTM|LDTM)"),
        ("UTCBAR (tcgen05.commit)", r"^UTCBAR"),
        ("LDGSTS (cp.async)", r"^LDGSTS"),
        ("UBLKCP / UTMALDG (bulk / TMA)", r"^(UBLKCP|UTMALDG)"),
        ("LDS.128", r"^LDS\.128"),
        ("BAR.SYNC", r"^BAR"),
        ("ACQBULK (griddepcontrol.wait)", r"^ACQBULK")]

out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
fam = collections.OrderedDict()
cur = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"^void ", "", name)
        name = re.sub(r"<.*", "", name.split("(")[0]).split("::")[-1]
        cur = fam.setdefault(name, {"n": 0, "c": collections.Counter()})
        cur["n"] += 1
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m and cur is not None:
        op = m.group(1)
        for title, pat in COLS:
            if re.search(pat, op):
                cur["c"][title] += 1
print("| kernel family | instantiations | " + " | ".join(t for t, _ in COLS) + " |")
print("|---|---|" + "---|" * len(COLS))
for name, d in fam.items():
    print("| `%s` | %d | " % (name, d["n"]) + " | ".join(str(d["c"][t]) for t, _ in COLS) + " |")
