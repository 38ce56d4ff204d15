"""Per-kernel SASS digest of librldl_b200.so (cuobjdump -sass): instruction counts that evidence the design claims —
UBLKCP (cp.async.bulk / TMA engine), SYNCS (mbarrier), DFMA/DADD/DMUL (FP64 pipe), LDS/STS, LDG/STG, SHFL, BAR,
no DMMA / HMMA / tensor-core instructions.  python tools/sass_digest.py > profiles/r02_sass_digest.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "osqp-recursive-ldl_b200", "lib", "librldl_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pats = ["UBLKCP", "SYNCS", "DFMA", "DADD", "DMUL", "MUFU", "LDS", "STS", "LDG", "STG", "SHFL", "BAR", "WARPSYNC", "ATOM", "RED",
        "DMMA", "HMMA", "UTCHMMA", "UTMALDG"]
cur, cnt = None, collections.OrderedDict()
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur).replace("void rldl::", "").replace("rldl::", "")
        cnt[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        cnt[cur]["total"] += 1
        for p in pats:
            if op == p or op.startswith(p + "."):
                cnt[cur][p] += 1
print("SASS digest of osqp-recursive-ldl_b200/lib/librldl_b200.so (sm_100a), instructions per kernel")
print(f"{'kernel':70s} {'total':>7s} " + " ".join(f"{p:>8s}" for p in pats))
for k, c in cnt.items():
    print(f"{k[:70]:70s} {c['total']:7d} " + " ".join(f"{c[p]:8d}" for p in pats))
