import re, subprocess, collections
out = subprocess.run(["cuobjdump", "-sass", "dalle_b200/ops/libdalle_b200.so"], capture_output=True, text=True).stdout
cur, stats = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()[:110]
        stats[cur] = collections.Counter()
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        full = m.group(1)
        op = full.split(".")[0]
        if op == "UTCHMMA" and ".2CTA" in full:
            op = "UTCHMMA.2CTA"
        if op in ("UTCHMMA", "UTCHMMA.2CTA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "UTCATOMSWS", "SYNCS", "UBLKCP", "FFMA2", "FMUL2", "FADD2", "RED", "MUFU"):
            stats[cur][op] += 1
print("# SASS evidence (cuobjdump -sass dalle_b200/ops/libdalle_b200.so, sm_100a): tensor-core / TMA mnemonics per kernel")
print("# UTC*MMA = tcgen05.mma (.2CTA = cta_group::2), LDTM = tcgen05.ld, UTMALDG/UTMASTG = cp.async.bulk.tensor load/store,")
print("# UBLKCP = cp.async.bulk, SYNCS = mbarrier, FFMA2/FMUL2/FADD2 = packed fp32x2 math, RED = red.global.add\n")
for k, c in stats.items():
    if c:
        print(k)
        print("    " + ", ".join(f"{op}x{n}" for op, n in sorted(c.items())))
