"""Per-kernel counts of the SASS instructions that show how libncdrift.so moves bytes on sm_100a.

    python profiles/sass_grep.py > profiles/sass_grep.txt

UBLKCP = cp.async.bulk (the TMA engine's 1-D bulk copy), LDGSTS = cp.async (16-byte global-to-shared copies), SYNCS = mbarrier arrive / try_wait, REDUX = warp
reductions in hardware, LDG.E.128 = 16-byte global loads, UTMALDG / UTCMMA / tcgen05 would be tensor-map TMA and
5th-generation tensor cores (none expected: nothing here is a contraction).  Built from the library in the tree.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "nexus_configuration_controller_b200", "lib", "libncdrift.so")
# label -> regex over one SASS line
PATTERNS = collections.OrderedDict([
    ("UBLKCP", r"\bUBLKCP"), ("LDGSTS", r"\bLDGSTS"), ("SYNCS.ARRIVE", r"\bSYNCS\.ARRIVE"), ("SYNCS.TRYWAIT", r"\bSYNCS\.PHASECHK"), ("REDUX", r"\bREDUX"),
    ("LDG.128", r"\bLDG\.E[.A-Z0-9]*\.128"), ("STG.128", r"\bSTG\.E[.A-Z0-9]*\.128"), ("LDS.128", r"\bLDS[.A-Z0-9]*\.128"),
    ("ATOMG/RED", r"\b(ATOMG|RED)\b|\b(ATOMG|RED)\."), ("SHFL", r"\bSHFL"), ("VOTE", r"\bVOTE"),
    ("UTMALDG", r"\bUTMALDG"), ("UTCMMA", r"\bUTC[A-Z]*MMA"), ("HMMA", r"\b[HIDQ]MMA"),
])


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", out)))
    counts, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = name.replace("(anonymous namespace)::", "").split("(")[0]
            cur = counts.setdefault(name, collections.Counter())
            continue
        if cur is None or "/*" not in line:
            continue
        for p, rx in PATTERNS.items():
            if re.search(rx, line):
                cur[p] += 1
        cur["instructions"] += 1 if re.search(r"/\*[0-9a-f]{4}\*/", line) else 0
    print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)}   cubin arch: {', '.join(arch)}")
    used = [p for p in PATTERNS if any(c[p] for c in counts.values())]
    print(f"{'kernel':44s} {'instr':>6s} " + " ".join(f"{p:>13s}" for p in used))
    for name, c in counts.items():
        print(f"{name[:44]:44s} {c['instructions']:6d} " + " ".join(f"{c[p]:13d}" for p in used))
    absent = [p for p in PATTERNS if p not in used]
    print("# never emitted: " + ", ".join(absent))


if __name__ == "__main__":
    sys.exit(main())
