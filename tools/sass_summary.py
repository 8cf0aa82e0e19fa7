"""Counts the Blackwell-specific SASS mnemonics per kernel of the built objects (cuobjdump -sass):
UTCHMMA (tcgen05.mma), UTCBAR (tcgen05.commit), LDTM / STTM (tcgen05.ld / st), UTMALDG (TMA tensor load),
UBLKCP (cp.async.bulk), SYNCS (mbarrier), STG.*256 (32-byte stores).  usage: python tools/sass_summary.py > profiles/rNN_sass_summary.txt"""
import collections
import glob
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MN = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UBLKCP", "SYNCS", "STG.E.ENL2.256", "UTCATOMSWS", "MUFU"]
print("%-110s %8s " % ("kernel", "instrs") + " ".join("%8s" % m[:8] for m in MN))
for obj in sorted(glob.glob(os.path.join(ROOT, "rave_b200", "build", "*.o"))):
    out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    name, cnt, total = None, collections.Counter(), 0
    rows = []

    def flush():
        if name and (cnt["UTCHMMA"] or cnt["UBLKCP"] or cnt["UTMALDG"] or cnt["LDTM"]):
            rows.append((name, total, dict(cnt)))
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            flush()
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(rb::tc::TcParams, rb::tc::TcTmap\)", "", name).replace("void rb::tc::", "")
            cnt, total = collections.Counter(), 0
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            total += 1
            op = m.group(1)
            for k in MN:
                if op.startswith(k):
                    cnt[k] += 1
    flush()
    for name, total, c in rows:
        print("%-110s %8d " % ((os.path.basename(obj) + ": " + name)[:110], total) + " ".join("%8d" % c.get(m, 0) for m in MN))
