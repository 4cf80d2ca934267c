"""Per-kernel counts of the SASS mnemonics that prove tcgen05 / TMEM / TMA use (profiles/rNN_sass_markers.txt).
Usage: python tools/sass_markers.py [path/to/liblrscore.so] > profiles/r02_sass_markers.txt"""
import os
import re
import subprocess
import sys

so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                        "mlapi_b200", "liblrscore.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
COLS = [("UTCHMMA", r"\bUTCHMMA\b(?!\.2CTA)"), (".2CTA", r"\bUTCHMMA\.2CTA"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"),
        ("UTMALDG", r"\bUTMALDG"), ("MCAST", r"UTMALDG[.\w]*MULTICAST"), ("UBLKCP", r"\bUBLKCP"), ("UTCBAR", r"\bUTCBAR"),
        ("LDGSTS", r"\bLDGSTS"), ("HMMA", r"\bHMMA")]
print("SASS markers per kernel of mlapi_b200/liblrscore.so (tools/sass_markers.py: cuobjdump -sass, count of mnemonic occurrences), round 2.")
print("UTCHMMA = tcgen05.mma kind::f16/tf32 (.2CTA = cta_group::2), LDTM / STTM = tcgen05.ld / .st, UTMALDG = TMA tensor load")
print("(.MULTICAST only in the opt-in W multicast of gemm_softmax_kernel: its tensor map covers the library's own copy of W on")
print("this GPU), UBLKCP = cp.async.bulk (1-D), UTCBAR = tcgen05.commit, LDGSTS = cp.async (rows TMA cannot address).  No HMMA")
print("(legacy mma.sync) anywhere.\n")
print(f"{'kernel':90s}" + "".join(f"{n:>9s}" for n, _ in COLS))
name, body, totals = None, [], [0] * len(COLS)


def flush():
    if name is None:
        return
    text = "\n".join(body)
    counts = [len(re.findall(rx, text)) for _, rx in COLS]
    for i, c in enumerate(counts):
        totals[i] += c
    print(f"{name[:90]:90s}" + "".join(f"{c:9d}" for c in counts))


for line in sass.split("\n"):
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        flush()
        name, body = m.group(1), []
    elif name is not None:
        body.append(line)
flush()
print(f"{'total':90s}" + "".join(f"{c:9d}" for c in totals))
