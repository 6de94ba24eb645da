"""Blackwell-native evidence: counts of the SASS mnemonics that prove tcgen05 / TMEM / TMA (B200_PROFILING.md) per kernel of
the in-tree library.    python tools/sass_table.py > profiles/r02_sass_table.txt"""
import re, subprocess, sys
from pathlib import Path
lib = Path(__file__).resolve().parents[1] / "yoho_b200" / "lib" / "libyoho_b200.so"
sass = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
WANT = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTCBAR", "UBLKCP", "LDGSTS", "FFMA2", "FADD2", "FMUL2", "HMMA", "MUFU.EX2", "MUFU.LG2", "SYNCS"]
cur, table = None, {}
for line in sass.splitlines():
    m = re.match(r"\s+Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", "-p", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
        table[cur] = dict.fromkeys(WANT, 0); table[cur]["instructions"] = 0
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1); table[cur]["instructions"] += 1
        for w in WANT:
            if w == "UTCHMMA.2CTA":
                if op.startswith("UTCHMMA") and ".2CTA" in op: table[cur][w] += 1
            elif w == "UTCHMMA":
                if op.startswith("UTCHMMA") and ".2CTA" not in op: table[cur][w] += 1
            elif op.startswith(w): table[cur][w] += 1
print(f"{lib.name}: SASS mnemonic counts per kernel (cuobjdump -sass); tcgen05.mma = UTCHMMA, tcgen05.ld/st = LDTM/STTM, TMA = UTMA*, cp.async = LDGSTS, packed fp32 = F*2")
cols = ["instructions"] + WANT
print(f"{'kernel':72s} " + " ".join(f"{c[:9]:>9s}" for c in cols))
for k, v in sorted(table.items()):
    if v["instructions"] < 40: continue
    print(f"{k[:72]:72s} " + " ".join(f"{v[c]:9d}" for c in cols))
