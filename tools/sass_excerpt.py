"""Writes profiles/<round>_sass_excerpt.md: the 128-bit loads, shared-memory broadcasts, warp votes / reductions and min-max
network instructions of the hot kernels, from `cuobjdump -sass` of the built library.  usage: python tools/sass_excerpt.py r02"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
txt = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "fast-lio-sam_gps_b200", "lib", "liblio_b200.so")], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
out = []


def grab(name_sub, label, pats, maxlines=60):
    for f in funcs:
        head = f.split("\n", 1)[0]
        if name_sub not in head:
            continue
        ops = [l for l in f.split("\n") if re.search(r"/\*[0-9a-f]{4}\*/", l)]
        counts = {}
        for l in ops:
            m = re.search(r"/\*[0-9a-f]{4}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
            if m and any(m.group(2).startswith(p) for p in pats):
                counts[m.group(2)] = counts.get(m.group(2), 0) + 1
        out.append(f"## {label}\n`{head.strip()}`\n")
        out.append(f"{len(ops)} SASS instructions; of the listed kinds (static counts): " + ", ".join(f"{k} x{v}" for k, v in sorted(counts.items(), key=lambda kv: -kv[1])) + "\n")
        sel = [l.rstrip() for l in ops if any(re.search(r"\s" + re.escape(p), l) for p in pats)]
        out.append("```\n" + "\n".join(sel[:maxlines]) + ("\n..." if len(sel) > maxlines else "") + "\n```\n")
        return


grab("k_knn_tileILb1ELb0", "k_knn_tile<XFORM, host loop> (IEKF search)",
     ["LDG.E.128", "LDS.128", "VIMNMX", "STG.E.128", "LDG.E.64", "ATOMG", "REDUX", "CREDUX", "VOTE", "LDG.E.CONSTANT"], 70)
grab("k_knn_leftILb1ELb0", "k_knn_left<XFORM, host loop> (per-query search)", ["LDG.E.128", "LDS", "STG.E.128"], 30)
grab("k_plane_reduceILi6ELb0", "k_plane_reduce<6 columns, host loop>", ["LDG.E.128", "DFMA", "LDS.64", "LDS.128", "STG.E.128"], 30)
grab("k_icp_nnILi0", "k_icp_nn<0> (ICP correspondence search)", ["LDG.E.128", "STG.E.128", "LDL", "STL"], 40)
hdr = ("# SASS excerpts (cuobjdump -sass of fast-lio-sam_gps_b200/lib/liblio_b200.so, sm_100a)\n\nWhat north_star asks to see: 128-bit coalesced "
       "global loads (`LDG.E.128`), 128-bit shared-memory broadcasts (`LDS.128`) in the candidate scan, and the branch-free `VIMNMX` "
       "insertion network of the tile search.  Written by tools/sass_excerpt.py.\n\n")
open(os.path.join(ROOT, "profiles", f"{tag}_sass_excerpt.md"), "w").write(hdr + "\n".join(out))
print("wrote", f"profiles/{tag}_sass_excerpt.md")
