"""Counts of the Blackwell-specific SASS instructions per kernel of libegpt_b200.so (cuobjdump -sass), so that the
tcgen05 / TMEM / TMA claims can be checked without rebuilding.  Usage: python tools/sass_summary.py > profiles/r02_sass_summary.md"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "eventgpt_b200", "lib", "libegpt_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
PAT = ["UTCHMMA.2CTA", "UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "SYNCS", "IDP.2A", "IDP.4A",
       "VIMNMX3.U16x2", "STS.U8", "LDS.U8", "REDUX", "MATCH.ANY", "MUFU.EX2", "MUFU.TANH", "RED.E", "ATOMS", "ATOMG", "HMMA", "BAR.SYNC", "BAR.RED"]
kern, rows = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0].replace("void ", "")
        rows[kern] = collections.Counter()
        continue
    if kern and re.search(r"/\*[0-9a-f]{4,5}\*/", line):
        rows[kern]["instructions"] += 1
        for p in PAT:
            if re.search(r"\b" + re.escape(p) + r"\b", line):
                rows[kern][p] += 1
                break
print("# Round 2 — Blackwell-specific SASS per kernel of `libegpt_b200.so` (`cuobjdump -sass`, sm_100a)\n")
print("`UTCHMMA` = tcgen05.mma (`.2CTA` = cta_group::2), `LDTM/STTM` = tcgen05.ld/st (TMEM), `UTMALDG/UTMASTG/UTMAREDG` = TMA tensor load / store / "
      "reduce, `UBLKCP` = cp.async.bulk (1-D bulk copy), `SYNCS` = mbarrier ops, `IDP` = dp2a/dp4a field extraction, `HMMA` = legacy mma.sync (none expected).\n")
cols = [p for p in PAT if any(r[p] for r in rows.values())]
print("| kernel | instr | " + " | ".join(cols) + " |\n|---|---:|" + "---:|" * len(cols))
for k, r in rows.items():
    if r["instructions"] < 50 and not any(r[p] for p in cols):
        continue
    print(f"| `{k}` | {r['instructions']} | " + " | ".join(str(r[p]) if r[p] else "" for p in cols) + " |")
