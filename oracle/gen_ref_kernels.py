#!/usr/bin/env python3
"""TEST INFRASTRUCTURE (oracle/): generates oracle/_ref/gen/*.inc from the reference tree where it lies.

Nothing produced here is committed (oracle/_ref/ is git-ignored).  Called by oracle/Makefile's `ref`
target, only when /root/reference exists (the build container).

  ref_kernels.inc     the UNMODIFIED OpenCL C text of the Lift kernels
                          example/{spmv,sssp,bfs,pr,scc}/kernel5.json:3   (glb-sdp, padded ELL, reversed order)
                          example/{spmv,sssp,bfs,pr,scc}/kernel6.json:3   (glb-sdp-rsa, ragged, forward order)
                      each wrapped in its own C++ namespace k_<alg>_<5|6> so that the ten KERNEL symbols and
                      their user functions (absadd, clmin, bool_and, int_min, ...) do not collide.  The text
                      compiles as C++ once `kernel`, `global`, `restrict`, get_global_id/size, fabs, min and
                      max are shimmed (oracle/ref_kernels_dev.cpp).
  ref_generators.inc  the initial-state vector generator classes of the apps, cut verbatim from
                          app/sssp.cpp:179-209, app/bfs.cpp:177-207, app/scc.cpp:179-195
                      (everything between the first `class ...Generator...` template and `int main`), each in a
                      namespace g_<alg>.
"""
import json
import os
import re
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "gen")
ALGS = ["spmv", "sssp", "bfs", "pr", "scc"]


def main():
    os.makedirs(OUT, exist_ok=True)
    with open(os.path.join(OUT, "ref_kernels.inc"), "w") as f:
        f.write("// GENERATED from %s/example/<alg>/kernel{5,6}.json (field \"source\"); do not commit\n" % REF)
        for alg in ALGS:
            for k in (5, 6):
                path = os.path.join(REF, "example", alg, "kernel%d.json" % k)
                j = json.load(open(path))
                want = ["MHeight", "MWidthC", "VLength"] if k == 5 else ["MHeight", "MWidthC"]
                assert j["paramVars"] == want, (path, j["paramVars"])
                f.write("\n// ---- %s  name=%s properties=%s\n" % (path, j["name"], json.dumps(j["properties"])))
                f.write("namespace k_%s_%d {\n%s\n}\n" % (alg, k, j["source"]))
    with open(os.path.join(OUT, "ref_generators.inc"), "w") as f:
        f.write("// GENERATED from %s/app/{sssp,bfs,scc}.cpp (generator classes); do not commit\n" % REF)
        for alg in ("sssp", "bfs", "scc"):
            src = open(os.path.join(REF, "app", alg + ".cpp")).read()
            end = src.index("int main(")
            m = re.search(r"template <typename T>\s*\nclass \w*Generator\w*", src)
            assert m and m.start() < end, alg
            f.write("\nnamespace g_%s {\n%s\n}\n" % (alg, src[m.start():end]))


if __name__ == "__main__":
    main()
