"""Blackwell evidence from the built library: per kernel, how many tcgen05 / TMA / TMEM instructions its SASS carries.
usage: python tools/sass_hist.py [path/to/libniti_b200.so] > profiles/sass_rN.txt
(tcgen05.mma -> UTC*MMA, tcgen05.ld/st -> LDTM/STTM, cp.async.bulk.tensor -> UTMALDG/UTMASTG/UTMAREDG, mbarrier -> SYNCS,
 tcgen05.commit -> UTCBAR, tcgen05.alloc -> UTCATOMSWS; /opt/skills/guides/B200_PROFILING.md)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = ("UTCIMMA", "UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "LDTM", "STTM", "UTCBAR", "UTCATOMSWS", "SYNCS",
        "UCGABAR", "HMMA", "IMMA", "LDGSTS", "IDP", "VIMNMX3", "ELECT", "ACQBULK", "REDG", "MEMBAR")


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "mandheling_dsp_training_b200", "libniti_b200.so")
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(anonymous namespace\)::", "", name)
            cur = kernels.setdefault(name.split("(")[0][:70], collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Z0-9_]+)*)", line)
        if m and cur is not None:
            cur["_total"] += 1
            op = m.group(1)
            full = op + m.group(2)
            for w in WANT:
                if op == w or op.startswith(w):
                    key = w
                    if w in ("UTCIMMA", "UTMALDG", "UTCBAR") and (".2CTA" in full):
                        key += ".2CTA"
                    if w == "UTMALDG" and "IM2COL" in full:
                        key += ".IM2COL"
                    cur[key] += 1
    print(f"# cuobjdump -sass {os.path.relpath(lib, ROOT)}: instruction counts per kernel (static), sm_100a")
    for name, c in kernels.items():
        keys = [k for k in c if k != "_total"]
        if not keys:
            continue
        print(f"{name:72s} {c['_total']:6d} instr | " + ", ".join(f"{k} {c[k]}" for k in sorted(keys)))


if __name__ == "__main__":
    main()
