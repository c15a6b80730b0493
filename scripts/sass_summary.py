"""SASS evidence from the built kernel library (runs on the CPU box: cuobjdump only).

  profiles/sass_mnemonics_r2.txt    per kernel: counts of the mnemonics that prove the Blackwell paths (UTC*MMA = tcgen05.mma,
                                    LDTM = tcgen05.ld, UTMALDG / UTMASTG / UTMAREDG = TMA tensor load / store / reduce, UBLKCP =
                                    cp.async.bulk, SYNCS = mbarrier, multimem = switch reductions, system-scope loads / stores)
  profiles/sass_comm_optim_r2.txt   full SASS listings of the collective and optimizer kernels
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "determined_b200", "_lib", "libdet_kernels.so")
KEEP = re.compile(r"^(UTC\w*MMA|UTCBAR|UTCATOMSWS|LDTM|STTM|UTMALDG|UTMASTG|UTMAREDG|UTMAPF|UBLKCP|UBLKRED|SYNCS|HMMA|MULTIMEM|"
                  r"LDG|STG|RED|REDG|ATOMG|ATOM|MEMBAR|ERRBAR|FENCE|CCTL|BAR|UCGABAR\w*|ACQBULK|LDS|STS|R2UR)")
FULL = ("fused_allreduce", "p2p_broadcast", "metric_allreduce", "fused_opt_kernel", "sumsq_finite", "noalloc")


def main() -> None:
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts = collections.defaultdict(collections.Counter)
    listings = collections.defaultdict(list)
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
            cur = cur.replace("void ", "")
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", line)
        if cur and m:
            op = m.group(1)
            if any(k in cur for k in FULL):
                listings[cur].append(line.split("/*", 2)[1].split("*/", 1)[1].rsplit("/*", 1)[0].rstrip() if "/*" in line else line)
            if KEEP.match(op):
                # keep scope / width qualifiers that matter as evidence, drop register-specific tails
                key = ".".join(op.split(".")[:4])
                counts[cur][key] += 1
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    with open(os.path.join(ROOT, "profiles", "sass_mnemonics_r2.txt"), "w") as f:
        for k in sorted(counts):
            interesting = {o: n for o, n in counts[k].items() if not re.match(r"^(LDS|STS|BAR|LDG\.E$|STG\.E$)", o)}
            if not interesting:
                continue
            f.write(k + "\n")
            for o, n in sorted(interesting.items()):
                f.write(f"    {n:5d}  {o}\n")
    with open(os.path.join(ROOT, "profiles", "sass_comm_optim_r2.txt"), "w") as f:
        for k in sorted(listings):
            f.write(f"==== {k} ({len(listings[k])} instructions)\n")
            f.write("\n".join(listings[k]) + "\n\n")
    print("kernels:", len(counts), "listings:", len(listings))


if __name__ == "__main__":
    sys.exit(main())
