"""Per-kernel counts of the Blackwell-specific SASS instructions in libddnn.so (cuobjdump -sass): tcgen05 MMAs
(UTCHMMA / UTCQMMA..., .2CTA for cta_group::2), TMA loads (UTMALDG), TMEM loads (LDTM), tcgen05 commits (UTCBAR),
legacy tensor-core MMAs (HMMA), cluster barriers / distributed shared memory.  Writes profiles/sass_summary.txt.
    python tools/sass_summary.py"""
import collections
import os
import re
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(REPO, "docker-dist-nn_b200", "libddnn.so")
PATTERNS = ["UTCHMMA.2CTA", "UTCHMMA", "UTMALDG.2D.2CTA", "UTMALDG", "LDTM", "UTCBAR.2CTA.MULTICAST", "UTCBAR", "UTCATOMSWS",
            "HMMA.1688.F32.TF32", "HMMA.16816.F32.BF16", "UCGABAR", "SYNCS", "LDGSTS", "ACQBULK", "MAPA", "LDS.128", "FFMA"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
    counts = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
            cur = re.sub(r"\(CUtensorMap_st.*", "", cur)
            cur = re.sub(r"ddnn::\(anonymous namespace\)::", "", cur).replace("void ", "")
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        for p in PATTERNS:
            if re.search(r"\b" + re.escape(p) + r"(\b|\.)", line):
                counts[cur][p] += 1
                break
    lines = ["# SASS instruction counts per kernel of docker-dist-nn_b200/libddnn.so (sm_100a), from cuobjdump -sass.",
             "# Kernels that use the tensor cores / TMA / TMEM are listed; totals over all kernels at the end.", ""]
    total = collections.Counter()
    for k, c in counts.items():
        total.update(c)
        if any(c[p] for p in PATTERNS if p.startswith(("UTC", "UTMA", "LDTM", "HMMA"))):
            lines.append(f"{k[:150]}")
            lines.append("    " + ", ".join(f"{p}={c[p]}" for p in PATTERNS if c[p]))
    lines += ["", f"kernels in the library: {len(counts)}", "totals: " + ", ".join(f"{p}={total[p]}" for p in PATTERNS if total[p])]
    path = os.path.join(REPO, "profiles", "sass_summary.txt")
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    print("\n".join(lines[-2:]))
    print("wrote", path)


if __name__ == "__main__":
    sys.exit(main())
