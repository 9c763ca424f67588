"""Counts of the SASS mnemonics that prove the Blackwell / NVLink-specific paths, per kernel of libgsx.so.
usage: python profiles/sass_summary.py > profiles/r2_sass_summary.txt"""
import collections
import re
import subprocess

out = subprocess.run(["cuobjdump", "-sass", "3dgs_b200/libgsx.so"], capture_output=True, text=True).stdout
pat = {"UBLKCP (TMA bulk copy)": r"\bUBLKCP", "UBLKRED (TMA bulk reduce-add)": r"\bUBLKRED", "LDGSTS (cp.async)": r"\bLDGSTS",
       "LDGMC.*ADD (multimem.ld_reduce)": r"\bLDGMC", "REDG.E.ADD.F32x4 (red.global.add.v4.f32)": r"REDG?\.E\.ADD\.F32x4",
       "REDG.E.ADD.F32 (scalar RED)": r"REDG?\.E\.ADD\.F32\.", "FFMA2/FMUL2/FADD2 (packed fp32x2)": r"\b(FFMA2|FMUL2|FADD2)\b",
       "MUFU.EX2": r"MUFU\.EX2", "MUFU.RCP": r"MUFU\.RCP", "SHFL": r"\bSHFL", "ATOMS.CAST.SPIN (shared float atomic loop)": r"ATOMS\.CAST",
       "STG...SYS on multicast (multimem.st)": r"STG\.E(\.128)?\.STRONG\.SYS", "SYNCS / mbarrier": r"\bSYNCS"}
cur, counts = None, collections.OrderedDict()
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(anonymous namespace\)::|gsx::", "", cur)
        cur = re.sub(r"\(.*", "", cur)[:70]
        counts[cur] = collections.Counter()
        continue
    if cur:
        for k, p in pat.items():
            if re.search(p, ln):
                counts[cur][k] += 1
print("# cuobjdump -sass 3dgs_b200/libgsx.so: mnemonic counts per kernel (static instruction counts)")
for k, c in counts.items():
    if c:
        print(f"{k}\n    " + ", ".join(f"{n} x{v}" for n, v in c.items()))
