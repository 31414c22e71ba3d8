#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove the tcgen05 / TMEM / TMA path (B200_PROFILING.md's list), from
`cuobjdump -sass` of the shipped library.  usage: python profiles/tools/sass_ops.py [lib.so] > profiles/rNN_sass_ops.txt"""
import collections
import re
import subprocess
import sys

LIB = sys.argv[1] if len(sys.argv) > 1 else "scaleprotoseg_b200/libscaleproto_sm100.so"
OPS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTMAPF",
       "SYNCS", "MUFU.EX2", "MUFU.LG2", "MUFU.RCP", "HMMA", "LDGSTS", "ACQBULK", "BAR.SYNC"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    cur, counts, arch = None, collections.OrderedDict(), set()
    for line in sass.split("\n"):
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        m = re.match(r"\s*arch = (\S+)", line)
        if m:
            arch.add(m.group(1))
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        counts[cur]["_total"] += 1
        for o in OPS:
            if op == o or op.startswith(o + ".") or (("." in o) and op.startswith(o)):
                counts[cur][o] += 1
    names = demangle(list(counts))
    print(f"# {LIB}: cuobjdump -sass, arch {sorted(arch)}; static instruction counts per kernel (not executed counts)")
    print("# UTCHMMA = tcgen05.mma kind::f16, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UTMALDG = cp.async.bulk.tensor (TMA tile"
          " load), UBLKCP = cp.async.bulk (1-D bulk copy), SYNCS = mbarrier ops")
    agg = collections.OrderedDict()
    for k, c in counts.items():
        short = re.sub(r"\(.*", "", names[k].replace("(anonymous namespace)::", "")).replace("void ", "")
        base = re.sub(r"<.*", "", short)
        a = agg.setdefault(base, [0, collections.Counter(), []])
        a[0] += 1
        a[2].append((short, c))
    print("\n## per kernel family (instantiations; range of counts over the instantiations)")
    for base, (n, _, inst) in agg.items():
        parts = []
        for o in OPS:
            v = [c[o] for _, c in inst]
            if max(v) > 0:
                parts.append(f"{o} {min(v)}" + (f"-{max(v)}" if max(v) != min(v) else ""))
        tot = [c["_total"] for _, c in inst]
        print(f"{base:32s} x{n:<3d} instr {min(tot)}-{max(tot)}  | " + ", ".join(parts))
    print("\n## every instantiation")
    for base, (n, _, inst) in agg.items():
        for short, c in inst:
            parts = [f"{o} {c[o]}" for o in OPS if c[o] > 0]
            print(f"{short[:110]:110s} | instr {c['_total']} | " + ", ".join(parts))


if __name__ == "__main__":
    main()
