"""SASS opcode histogram per kernel of the built library (cuobjdump -sass) -> profiles/sass_opcodes_r2.txt."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "enarf-gan_b200", "libenarf_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
per, fn = collections.defaultdict(collections.Counter), None
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and fn:
        per[fn][m.group(1)] += 1
keys = ["HMMA", "UTCHMMA", "UTCMMA", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "LDGSTS", "RED", "REDG", "ATOMG", "ATOMS",
        "LDG", "STG", "LDS", "STS", "BAR", "SHFL", "FFMA", "FMUL", "FADD", "MUFU", "SYNCS"]
demangle = subprocess.run(["c++filt"], input="\n".join(per), capture_output=True, text=True).stdout.splitlines()
out = ["SASS opcode histogram of enarf-gan_b200/libenarf_b200.so (cuobjdump -sass, sm_100a), round 2 build.",
       "Per kernel: total instructions, then the opcodes that tell which hardware paths it uses (HMMA = mma.sync tensor",
       "instructions; UTC*MMA / LDTM / STTM = tcgen05 + TMEM; UBLKCP / UTMA* = TMA; LDGSTS = cp.async; RED = reductions).", ""]
for f, name in sorted(zip(per, demangle), key=lambda t: -sum(per[t[0]].values())):
    name = re.sub(r"\(.*", "", name)[:64]
    out.append(f"{name:64s} {sum(per[f].values()):7d}  " + "  ".join(f"{k} {per[f][k]}" for k in keys if per[f][k]))
open(os.path.join(ROOT, "profiles", "sass_opcodes_r2.txt"), "w").write("\n".join(out) + "\n")
print("\n".join(out))
