#!/usr/bin/env python
"""Instruction census of libcmf_b200.so from `cuobjdump -sass`: per kernel, how many tcgen05 MMA
(UTCHMMA / UTCQMMA ...), tensor-memory load/store (LDTM / STTM), TMA (UTMALDG / UTMASTG), bulk-copy
and mbarrier (SYNCS) instructions the SASS holds.  Evidence that the tensor-core kernels are tcgen05 +
TMA code and not recompiled mma.sync.   python scripts/sass_census.py > profiles/rNN_sass_census.txt"""
import collections, os, re, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cmf_b200", "lib", "libcmf_b200.so")
PAT = ["UTCHMMA", "UTCQMMA", "UTCIMMA", "UTCOMMA", "LDTM", "STTM", "UTCCP", "UTCBAR", "UTMALDG", "UTMASTG", "UBLKCP",
       "LDGSTS", "SYNCS", "HMMA", "IMMA", "FADD2", "FFMA2", "FMUL2"]


def main():
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    out = subprocess.run([cuobjdump, "-sass", LIB], capture_output=True, text=True).stdout
    demangle = shutil.which("cu++filt") or "/usr/local/cuda/bin/cu++filt"
    counts, total, cur = collections.OrderedDict(), {}, None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur], total[cur] = collections.Counter(), 0
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if cur and m:
            total[cur] += 1
            op = m.group(1).split(".")[0]
            for p in PAT:
                if op == p or (p in ("UTCHMMA", "UTCQMMA", "UTCIMMA", "UTCOMMA") and op.startswith(p)):
                    counts[cur][p] += 1
    names = list(counts)
    dem = subprocess.run([demangle] + names, capture_output=True, text=True).stdout.splitlines() if names else []
    print(f"# {os.path.relpath(LIB, ROOT)}: {len(names)} kernels (sm_100a SASS)")
    print(f"# {'kernel':70s} {'instrs':>7s}  " + " ".join(f"{p:>7s}" for p in PAT))
    for n, d in zip(names, dem):
        short = re.sub(r"\((?:cmf|CUtensorMap|int|float|long|unsigned|const).*", "", d).replace("cmf::", "").replace("(anonymous namespace)::", "")
        short = re.sub(r"^void ", "", short)
        print(f"{short[:110]:110s} {total[n]:7d}  " + " ".join(f"{counts[n][p]:7d}" for p in PAT))


if __name__ == "__main__":
    main()
