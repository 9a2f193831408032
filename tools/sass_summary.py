"""Count the Blackwell-specific SASS mnemonics per kernel of libdiner_b200.so (cuobjdump -sass), the evidence that the
MLP kernels are genuine tcgen05 / TMEM / bulk-copy-engine code:

    python tools/sass_summary.py > profiles/r2_sass_summary.md

UTCHMMA = tcgen05.mma (.2CTA = cta_group::2), LDTM / STTM = tcgen05.ld / st (TMEM), UTCBAR = tcgen05.commit,
UTMALDG = cp.async.bulk.tensor (tensor-map TMA), UBLKCP = cp.async.bulk (TMA engine, non-tensor), SYNCS = mbarrier ops, FFMA2 / FMUL2 = packed f32x2 math,
REDG / ATOMG = global reductions / atomics, CCTL.IVALL = L1 invalidate (cluster-scope acquire).
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "diner_b200", "libdiner_b200.so")
KEYS = ["UTCHMMA.2CTA", "UTCHMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UBLKPF", "SYNCS", "FFMA2", "FMUL2", "REDG", "RED.",
        "ATOMG", "CCTL.IVALL", "UTMALDG", "LDG", "STG", "LDS", "STS", "MUFU", "SHFL"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True).stdout
    regs = {}
    for m in re.finditer(r"Function (\S+):\s*\n\s*REG:(\d+).*?SHARED:(\d+)", res):
        regs[m.group(1)] = (int(m.group(2)), int(m.group(3)))
    counts, order, cur = collections.defaultdict(collections.Counter), [], None
    total = collections.Counter()
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            order.append(cur)
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m or cur is None:
            continue
        op = m.group(1)
        total[cur] += 1
        for k in KEYS:
            if k == "UTCHMMA" and op.startswith("UTCHMMA.2CTA"):
                continue
            if op.startswith(k):
                counts[cur][k] += 1
    demangle = subprocess.run(["cu++filt"] + order, stdout=subprocess.PIPE, text=True).stdout.splitlines()
    names = dict(zip(order, demangle))
    print("# SASS summary of `diner_b200/libdiner_b200.so` (sm_100a), `python tools/sass_summary.py`\n")
    print("Per kernel: SASS instructions, registers, static shared memory, and the counts of the mnemonics that identify the "
          "Blackwell units in use.\n")
    cols = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTCBAR", "UTMALDG", "UBLKCP", "UBLKPF", "SYNCS", "FFMA2", "FMUL2", "REDG", "ATOMG",
            "CCTL.IVALL", "MUFU", "SHFL"]
    print("| kernel | instr | regs | " + " | ".join(cols) + " |")
    print("|---|---:|---:|" + "---:|" * len(cols))
    tot = collections.Counter()
    for f in order:
        short = re.sub(r"\([^()]*\)\s*$", "", names.get(f, f)).replace("dnr::", "").replace("void ", "")
        c = counts[f]
        print(f"| `{short}` | {total[f]} | {regs.get(f, ('', ''))[0]} | " + " | ".join(str(c[k]) if c[k] else "" for k in cols) + " |")
        tot.update(c)
    print("| **all kernels** | %d | | " % sum(total.values()) + " | ".join(str(tot[k]) if tot[k] else "" for k in cols) + " |")
    print("\n`UTMALDG` (tensor-map TMA loads, %d): the quad kernel's operand ring in its default mode (cta_group::2 completion on "
          "the even CTA's barrier); everything else moves pre-packed operands with plain bulk copies (`UBLKCP`).  No `HMMA` / "
          "`HGMMA` (mma.sync / wgmma) anywhere: %d." % (tot["UTMALDG"], len(re.findall(r"\bHG?MMA\b", sass))))


if __name__ == "__main__":
    main()
