"""Per-kernel register / shared-memory / spill summary of a ptxas -v log (csrc/build/*.ptxas.log)."""
import re
import subprocess
import sys

for path in sys.argv[1:]:
    txt = open(path).read()
    blocks = re.findall(r"Compiling entry function '(\S+)' for 'sm_100a'\n.*\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n.*Used (\d+) registers(.*)\n", txt)
    if not blocks:
        continue
    names = subprocess.run(["c++filt"] + [b[0] for b in blocks], capture_output=True, text=True).stdout.splitlines()
    for (m, st, ss, sl, r, rest), n in zip(blocks, names):
        n = re.sub(r"\(.*", "", n.replace("(anonymous namespace)::", "")).replace("void ", "").replace("lio::", "")
        sm = re.search(r"(\d+) bytes smem", rest)
        print(f"{n:50s} regs={r:>3s} smem={sm.group(1) if sm else 0:>6} stack={st} spill={ss}/{sl}")
