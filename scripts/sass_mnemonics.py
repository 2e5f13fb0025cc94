#!/usr/bin/env python
"""Tensor-core / TMA / TMEM instruction counts per kernel of the built library (cuobjdump -sass), the evidence that the
hot kernels are tcgen05 / TMA code and which ones are mma.sync.  Usage: python scripts/sass_mnemonics.py > profiles/..."""
import collections
import re
import subprocess
import sys

LIB = "distributed_inference_server_b200/libdis_engine.so"
WANT = ("UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "HMMA", "LDSM", "SYNCS", "UTCCP")
out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
counts, cur = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur:
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            op = m.group(1)
            for w in WANT:
                if op.startswith(w):
                    counts[cur][w] += 1
print("# cuobjdump -sass libdis_engine.so (sm_100a): tensor-core / TMA / TMEM instructions per kernel (kernels with none are left out)")
print("# UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG = cp.async.bulk.tensor (TMA load), UTCBAR = tcgen05.commit,")
print("# SYNCS = mbarrier ops, HMMA / LDSM = mma.sync / ldmatrix")
for fn, c in counts.items():
    if not any(c[w] for w in WANT if w != "SYNCS"):
        continue
    name = re.sub(r"\(.*", "", demangle(fn))
    print(f"{name:70s} " + " ".join(f"{w}={c[w]}" for w in WANT if c[w]))
