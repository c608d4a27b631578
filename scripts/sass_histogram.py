"""Opcode histogram per kernel of the in-tree library: `cuobjdump -sass hop-vae_b200/libhop_sm100.so`, mnemonics that
prove the Blackwell-native paths (UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG = TMA, UBLKCP = bulk copy).
    python scripts/sass_histogram.py > profiles/r02_sass_opcodes.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "hop-vae_b200", "libhop_sm100.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
KEY = re.compile(r"^(UTC\w*MMA|UTCBAR|UTCCP|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|SYNCS|MUFU|FFMA|HMMA|RED|ATOM|LDGSTS|BAR|FMNMX\w*|F2FP|UCGABAR)")
kern, hist = None, {}
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = m.group(1)
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        op = m.group(1)
        base = op.split(".")[0]
        if KEY.match(base):
            full = ".".join(op.split(".")[:2]) if base.startswith(("UTC", "UTMA", "MUFU")) else base
            hist[kern][full] += 1
        hist[kern]["(all instructions)"] += 1
print(f"# cuobjdump -sass {os.path.relpath(lib, ROOT)}: selected opcodes per kernel (demangle with c++filt)")
for k in sorted(hist, key=lambda k: -hist[k]["(all instructions)"]):
    c = hist[k]
    name = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip()
    name = name.replace("(anonymous namespace)::", "").replace("void ", "")
    name = re.sub(r"\((CUtensorMap_st|hop::|int|float|const|unsigned).*", "", name)      # drop the parameter list
    print(f"\n{name}")
    print("  " + ", ".join(f"{op} {n}" for op, n in sorted(c.items(), key=lambda t: (-t[1], t[0]))))
tot = collections.Counter()
for c in hist.values():
    tot.update(c)
print("\n# library totals")
print("  " + ", ".join(f"{op} {n}" for op, n in sorted(tot.items(), key=lambda t: (-t[1], t[0]))))
