"""Parse the reference's bundled OBJ with the rules of CoveragePathPlanningHeuristic::loadOBJFile
(cscpp/src/coverage_path_planning_heuristic.cpp:1355-1451) and commit it as tests/golden/etihad_mesh.npz.
Runs in the build container only (/root/reference does not exist on the GPU box)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/cscpp/mesh/etihad_nowheels_nointernal_scaled_new.obj"


def load_obj(path):
    verts, tris = [], []
    has_normal = has_texture = False
    with open(path, "rb") as f:
        for raw in f:
            tok = raw.decode("ascii", "replace").replace("\r", " ").replace("\n", " ").replace("\t", " ").split(" ")
            tok = [t for t in tok if t]
            if not tok or tok[0][0] == "#":
                continue
            if tok[0][0] == "v":
                if len(tok[0]) > 1 and tok[0][1] == "n":
                    has_normal = True
                elif len(tok[0]) > 1 and tok[0][1] == "t":
                    has_texture = True
                else:
                    verts.append([np.float32(float(tok[1])), np.float32(float(tok[2])), np.float32(float(tok[3]))])  # (float)atof
            elif tok[0][0] == "f":
                data = tok[1:]
                n = len(data)
                for t in range(n - 2):
                    if not has_texture and not has_normal:
                        idx = [int(data[0]) - 1, int(data[1]) - 1, int(data[2]) - 1]  # :1404-1406 (always the first three)
                    else:
                        idx = [int(data[0].split("/")[0]) - 1] + [int(data[t + i].split("/")[0]) - 1 for i in (1, 2)]  # atoi stops at '/'
                    tris.append(idx)
    return np.array(verts, np.float32), np.array(tris, np.int32)


if __name__ == "__main__":
    v, t = load_obj(SRC)
    out = os.path.join(ROOT, "tests", "golden", "etihad_mesh.npz")
    np.savez_compressed(out, vertices=v, triangles=t, source=os.path.basename(SRC))
    print("vertices", v.shape, "triangles", t.shape, "->", out, os.path.getsize(out), "bytes")
