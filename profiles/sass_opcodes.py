"""SASS opcode histogram per kernel of the built library (cuobjdump -sass), tensor-core / TMA / TMEM opcodes first.
    python profiles/sass_opcodes.py > profiles/r2_sass_opcodes.txt"""
import collections, os, re, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(root, "deconv_api_b200", "_lib", "libdcv.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
key = ("UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "LDTM", "STTM", "UTCBAR", "UTCCP", "UTCATOMSWS", "UBLKCP", "SYNCS", "HMMA", "FFMA", "STS", "LDS", "SHFL")
kern, hist, total = None, {}, collections.Counter()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kern = re.sub(r"\((?!anonymous namespace\)).*", "", kern)       # drop the argument list, keep "(anonymous namespace)::"
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        op = m.group(1)
        hist[kern][op] += 1
        total[op.split(".")[0]] += 1
print("library:", os.path.relpath(lib, root), "- opcode counts from `cuobjdump -sass` (static instruction counts)")
print("whole library:", {k: total[k] for k in key if total[k]})
for k in sorted(hist, key=lambda n: -sum(v for o, v in hist[n].items() if o.startswith(("UTC", "UTMA", "LDTM", "UBLKCP")))):
    h = hist[k]
    tc = {o: v for o, v in sorted(h.items()) if o.startswith(("UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "UTCBAR", "UBLKCP", "SYNCS.ARRIVE", "SYNCS.PHASECHK"))}
    base = collections.Counter()
    for o, v in h.items():
        base[o.split(".")[0]] += v
    rest = {o: base[o] for o in ("FFMA", "HMMA", "STS", "LDS", "SHFL", "LDG", "STG") if base[o]}
    print("\n%s\n  instructions %d | tensor/TMA/TMEM: %s\n  other: %s" % (k, sum(h.values()), tc or "-", rest))
