#!/usr/bin/env python
"""Full SASS of every translation unit (gzip, profiles/sass/r2_<unit>.sass.gz) plus a readable index:
per kernel, registers and the count of the instructions that prove the Blackwell path (UTCHMMA = tcgen05.mma,
UTMALDG = TMA load, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, LDGMC/STGMC... = multimem, UCGABAR = cluster
barrier, HMMA = legacy tensor path which must stay at zero).  Run after python -m kubeml_b200._native.build."""
import gzip, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "build", "obj")
OUT = os.path.join(ROOT, "profiles", "sass")
os.makedirs(OUT, exist_ok=True)
KEYS = ["UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "UTCBAR", "UTCATOMSWS", "UCGABAR", "LDGMC", "STGMC", "MULTIMEM", "REDG",
        "RED.", "ATOM", "MEMBAR", "HMMA", "LDGSTS", "SYNCS", "BAR.SYNC"]
index = ["# SASS index (cuobjdump -sass of build/obj/*.o, sm_100a)\n",
         "Full listings: `profiles/sass/r2_<unit>.sass.gz` (zcat to read).  Counts are static instruction counts.\n",
         "| unit | kernel | SASS lines | " + " | ".join(KEYS) + " |", "|---|---|---|" + "---|" * len(KEYS)]
for o in sorted(os.listdir(OBJ)):
    if not o.endswith(".cu.o"):
        continue
    unit = o[:-5]
    sass = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, o)], capture_output=True, text=True).stdout
    with gzip.open(os.path.join(OUT, f"r2_{unit}.sass.gz"), "wt") as f:
        f.write(sass)
    for blk in sass.split("Function : ")[1:]:
        name = blk.split("\n", 1)[0].strip()
        dem = subprocess.run(["cu++filt", name], capture_output=True, text=True).stdout.strip() or name
        dem = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", dem).replace("void ", "")
        depth = 0
        for i in range(len(dem) - 1, -1, -1):          # cut the trailing "(parameter list)"
            depth += dem[i] == ")"
            depth -= dem[i] == "("
            if depth == 0 and dem[i] == "(":
                dem = dem[:i]
                break
        dem = dem.replace("(int)", "").replace("(bool)", "")
        lines = [l for l in blk.splitlines() if re.search(r"/\*[0-9a-f]{4}\*/", l)]
        counts = [sum(1 for l in lines if k in l) for k in KEYS]
        index.append(f"| {unit} | `{dem[:90]}` | {len(lines)} | " + " | ".join(str(c) for c in counts) + " |")
open(os.path.join(OUT, "r2_INDEX.md"), "w").write("\n".join(index) + "\n")
print("\n".join(index[:12]))
