"""profiles/r02_sass.md: counts and first occurrences of the SASS mnemonics that identify each hardware path, per kernel of
libnumsb200.so (cuobjdump -sass).  Usage: python tools/sass_evidence.py > profiles/r02_sass.md"""
import collections, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "nums_b200", "libnumsb200.so")
text = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
KEYS = ("UTCIMMA", "UTCBAR", "LDTM", "UTMALDG", "UBLKCP", "SYNCS", "DMMA", "UTCATOMSWS", "PRMT")
WANT = ("k_oz_gemm", "k_oz_rows", "k_tri_gram", "k_gemm_ws")
print("# Round 2: SASS evidence (cuobjdump -sass nums_b200/libnumsb200.so, sm_100a; tools/sass_evidence.py)\n")
print("Counts of the mnemonics that identify each path, and the first occurrences with their addresses. `tcgen05.mma` appears as "
      "`UTCIMMA` (integer kind), `tcgen05.ld` as `LDTM`, `tcgen05.commit` as `UTCBAR`, TMA tensor copies as `UTMALDG`, bulk copies "
      "as `UBLKCP`, mbarrier operations as `SYNCS`, the FP64 tensor path as `DMMA`, the byte transposes of the digit split as `PRMT`.\n")
fn = None
found = collections.OrderedDict()
for line in text.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1) if any(w in m.group(1) for w in WANT) else None
        if fn:
            found[fn] = collections.OrderedDict()
        continue
    if fn is None:
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if not m:
        continue
    addr, ins = m.group(1), m.group(2).strip()
    body = re.sub(r"^@!?U?P\d+\s+", "", ins)
    op = body.split()[0]
    if any(op.startswith(k) for k in KEYS):
        e = found[fn].setdefault(op, [0, addr, ins])
        e[0] += 1
for fn, ops in found.items():
    if not ops:
        continue
    print("\n## `%s`\n" % fn)
    print("| mnemonic | count | first at | instruction |\n|---|---|---|---|")
    for op, (c, addr, ins) in sorted(ops.items(), key=lambda kv: -kv[1][0]):
        print("| `%s` | %d | 0x%s | `%s` |" % (op, c, addr, ins))
