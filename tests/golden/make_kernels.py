"""Writes tests/golden/kernels/: kernel configuration files in the reference's JSON schema
(src/kernel_config.cpp:14-95) for the six Lift schedule families x four semirings,
plus two runfiles.  These are this repo's own fixtures: "source" carries only the user-function names that
identify the semiring (the CUDA build selects a precompiled kernel from them and never compiles the text).
The reference's real JSONs (example/<alg>/kernel*.json) parse the same way; tests use them when
/root/reference is present."""
import json
import os

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "kernels")
SEMIRINGS = {
    "spmv": "float add(float x, float y){ return x+y; }\nfloat mult(float l, float r){ return l*r; }\n",
    "sssp": "float clmin(float a, float b){ return fabs(a)<fabs(b)?fabs(a):fabs(b); }\n"
            "float absadd(float a, float b){ return fabs(a)+fabs(b); }\n",
    "bfs": "int bool_or(int a, int b){ return (a!=0)||(b!=0); }\nint bool_and(int a, int b){ return (a!=0)&&(b!=0); }\n",
    "scc": "int int_max(int a, int b){ return a>b?a:b; }\nint int_min(int a, int b){ return a<b?a:b; }\n",
}
ELL3 = ["MHeight", "MWidthC", "VLength"]
SCHEDULES = {
    "glb-sdp": ({"outerMap": "glb", "dotProduct": "sdp"}, ELL3, ["(4*v_MHeight_2*v_MWidthC_1)"]),
    "glb-sdp-rsa": ({"outerMap": "glb", "arrayType": "ragged", "dotProduct": "sdp"}, ELL3[:2], ["(4*v_MHeight_2)"]),
    "swrg-slcl-pmdp": ({"outerMap": "swrg", "innerMap": "slcl", "dotProduct": "pmdp"}, ELL3,
                       ["(4*v_MHeight_2*v_MWidthC_1)"]),
    "swrg-slcl-sdp-chunk-128": ({"outerMap": "swrg", "innerMap": "slcl", "chunkSize": "128", "dotProduct": "sdp"},
                                ELL3, ["(4*v_MHeight_2*v_MWidthC_1)"]),
    "awrg-alcl-alcl-edp-split-8": ({"outerMap": "awrg", "innerMap": "alcl", "innerMap2": "alcl", "splitSize": "8",
                                    "dotProduct": "edp"}, ELL3, ["4"]),
    "awrg-alcl-fdp-chunk-rsa-8": ({"outerMap": "awrg", "innerMap": "alcl", "chunkSize": "8", "arrayType": "ragged",
                                   "dotProduct": "fused"}, ELL3[:2], ["(4*v_MHeight_2)", "4"]),
}


def arg(v, space, size):
    return {"variable": v, "addressSpace": space, "size": size}


def main():
    for alg, src in SEMIRINGS.items():
        os.makedirs(os.path.join(OUT, alg), exist_ok=True)
        for name, (props, param_vars, temps) in SCHEDULES.items():
            j = {"name": name, "source": src + "kernel void KERNEL(){}\n", "properties": props,
                 "inputArgs": [arg("v__1", "global", "?"), arg("v__2", "global", "?"),
                               arg("v__3", "global", "(4*v_VLength_3)"), arg("v__4", "global", "(4*v_MHeight_2)"),
                               arg("v__5", "private", "4"), arg("v__6", "private", "4")],
                 "tempGlobals": [arg(f"v__t{i}", "global", t) for i, t in enumerate(temps)],
                 "outputArg": arg("v__o", "global", "(4*v_MHeight_2)"),
                 "tempLocals": [], "paramVars": param_vars, "outputSize": "(4*v_MHeight_2)"}
            with open(os.path.join(OUT, alg, name + ".json"), "w") as f:
                json.dump(j, f, indent=2)
    with open(os.path.join(OUT, "runfile.csv"), "w") as f:
        f.write("1280,1,1,1,1,1,\n")
    with open(os.path.join(OUT, "runfile2.csv"), "w") as f:
        f.write("16384,1,1,128,1,1,\n65536,1,1,256,1,1,\n")


if __name__ == "__main__":
    main()
