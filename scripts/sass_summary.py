"""profiles/sass_summary.txt: instruction counts per kernel of the built library (cuobjdump -sass), the mnemonics that
prove which hardware path a kernel takes.  Runs where the library was built (no GPU needed).

    python scripts/sass_summary.py > profiles/sass_summary.txt
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

lib = os.path.join(ROOT, g.PKG_DIRNAME if hasattr(g, "PKG_DIRNAME") else
                   "vt-cluster--parallel-and-distributed-programming-of-machine-learning-models_b200", "liblora_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True,
                       text=True, check=True).stdout.splitlines()
ops = ["UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "UTCBAR", "HMMA", "LDGSTS", "FFMA2"]
print("# cuobjdump -sass liblora_b200.so (sm_100a), instruction counts per kernel; tcgen05.mma = UTCHMMA, TMA = UTMALDG/UTMASTG,")
print("# tcgen05.ld = LDTM, tcgen05.commit = UTCBAR, mma.sync = HMMA, cp.async = LDGSTS; written by scripts/sass_summary.py")
print("kernel," + ",".join(ops))
for name, body in zip(names, re.split(r"Function : \S+", sass)[1:]):
    short = re.sub(r"\(.*", "", name.replace("(anonymous namespace)", "{anon}"))
    counts = [len(re.findall(r"\b" + op + r"\b", body)) for op in ops]
    print('"' + short + '",' + ",".join(str(c) for c in counts))
