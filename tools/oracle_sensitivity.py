"""How much do the UNPINNED third-party / libm choices move the reference's results?

The reference's own sources are compiled (oracle/Makefile) once per variant of the stand-in choices and run over every
fixture viewpoint of config 1 (2 x 1 225 poses of cscpp/txt, 2 x 17 583 poses of sspp/resources) and a fine-grid
synthetic structure; each variant is compared with the canonical build (the one the oracle and the CUDA path equal bit
for bit).  Output: profiles/r02_oracle_sensitivity.txt.

Variants
  dot4_seq    Eigen 4-wide dot / squaredNorm reduced ((a0+a1)+a2)+a3     (no vectorisation)
  dot4_hadd   (a0+a1)+(a2+a3)                                            (SSE3 haddps, -msse3 / -march=native builds)
  dot4_fma    fused multiply-add chain                                    (-mfma builds with contraction)
  dot3_seq    3-wide dot reduced (a0+a1)+a2 instead of a0+(a1+a2)
  tanf        tan() on the float argument resolved to the float overload  (frustum_culling.cpp:91-95)
  tan_double  ... forced to the double function
  sqrtf       sqrt() of the float32 squared distance in float             (voxel_grid_occlusion_estimation.cpp:560)
  sqrt_double ... forced to the double function
  math_h      <math.h> visible in the reference's translation units: all unqualified libm calls on floats go float
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from oracle import ref as R  # noqa: E402

SENS = "-include ../compat/sens/force_float_libm.h"
VARIANTS = [
    ("dot4_seq", "-DASSCPP_EIGEN_DOT4=1"),
    ("dot4_hadd", "-DASSCPP_EIGEN_DOT4=2"),
    ("dot4_fma", "-DASSCPP_EIGEN_DOT4=3"),
    ("dot3_seq", "-DASSCPP_EIGEN_DOT3=1"),
    ("tanf", SENS + " -DASSCPP_SENS_TANF"),
    ("tan_double", SENS + " -DASSCPP_SENS_TAN_DOUBLE"),
    ("sqrtf", SENS + " -DASSCPP_SENS_SQRTF"),
    ("sqrt_double", SENS + " -DASSCPP_SENS_SQRT_DOUBLE"),
    # <math.h> included before the reference's headers: libstdc++'s wrapper puts std::tan(float) / std::sqrt(float) /
    # std::floor(float) ... into the global namespace, so EVERY unqualified libm call on a float picks the float overload
    ("math_h", "-include math.h"),
]
NT = os.cpu_count() or 1


def workloads():
    g = os.path.join(ROOT, "tests", "golden")
    xyz = np.load(os.path.join(g, "etihad_model.npz"))["xyz"]
    a = np.load(os.path.join(g, "etihad_searchspace_2p5m.npz"))
    b = np.load(os.path.join(g, "etihad_searchspace_1p5m.npz"))
    poses = np.concatenate([a["cam0"], a["cam1"], b["cam0"], b["cam1"]])
    out = [("config 1: etihad 120 554 points, leaf 0.5 m, %d fixture poses" % len(poses), xyz, 0.0, poses)]
    import ctypes
    L = ctypes.CDLL(os.path.join(ROOT, "asscpp_b200", "lib", "libasscpp_host.so"))
    L.asscpp_synth_structure_v1.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int64, ctypes.c_void_p]
    L.asscpp_synth_views_v1.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int64, ctypes.c_void_p]
    sx = np.empty((1_000_000, 3), np.float32)
    sv = np.empty((2000, 7), np.float64)
    L.asscpp_synth_structure_v1(0xA55C0001, 0xA55C0002, len(sx), sx.ctypes.data)
    L.asscpp_synth_views_v1(0xA55C0001, 0xA55C0003, len(sv), sv.ctypes.data)
    out.append(("synthetic 1M points, leaf 0.05 m (the grid of configs 3/5), %d poses" % len(sv), sx, 0.05, sv))
    return out


def run(lib_path, xyz, leaf, poses):
    oc = R.RefOcclusionCulling(xyz=xyz, leaf=leaf, nthreads=NT, lib_path=lib_path)
    st = oc.eval_batch(poses)
    return oc, st


def main():
    lines = []

    def say(s=""):
        print(s)
        lines.append(s)

    say("# Sensitivity of the reference's results to the unpinned Eigen / libm choices (tools/oracle_sensitivity.py)")
    say("# every row: the reference's own sources (oracle/_ref) rebuilt with ONE choice changed, against the canonical build")
    say("# (canonical = what oracle/asscpp_oracle.cpp and the CUDA path reproduce bit for bit, tests/test_ref_pin.py)")
    R.build()
    wl = workloads()
    base = []
    for name, xyz, leaf, poses in wl:
        oc, st = run(R.LIB_PATH, xyz, leaf, poses)
        orc = O.Oracle(xyz, O.make_params(leaf=leaf if leaf > 0 else 0.5))
        base.append((oc, st, orc))
        say("")
        say("workload: %s" % name)
        say("  canonical: survivors %d, visible %d over all poses" % (st[:, 0].sum(), st[:, 1].sum()))
    for vname, defs in VARIANTS:
        libp = os.path.join(ROOT, "oracle", "_ref", "variants", "libasscpp_ref_%s.so" % vname)
        os.makedirs(os.path.dirname(libp), exist_ok=True)
        t = time.time()
        R.build(libp, defs=defs, objdir=os.path.join(ROOT, "oracle", "_ref", "variants", "obj_" + vname))
        say("")
        say("variant %-11s (%s)  [built in %.0f s]" % (vname, defs.replace(SENS, "-include force_float_libm.h"), time.time() - t))
        for (name, xyz, leaf, poses), (oc0, st0, orc) in zip(wl, base):
            oc, st = run(libp, xyz, leaf, poses)
            dv = np.nonzero((st[:, 0] != st0[:, 0]) | (st[:, 1] != st0[:, 1]) | (st[:, 2] != st0[:, 2]))[0]
            sbits = vbits = 0
            dpct = []
            for v in dv:
                s1, v1 = oc.extract(poses[v])
                s0, v0 = oc0.extract(poses[v])
                sbits += len(np.setxor1d(s0, s1))
                vbits += len(np.setxor1d(v0, v1))
                if leaf == 0.0:
                    dpct.append(abs(orc.coverage_percent_idx(v1) - orc.coverage_percent_idx(v0)))
            say("  %-62s views changed %5d / %d   survivor bits %6d / %d (%.2e)   visible bits %6d / %d (%.2e)%s" % (
                name[:62], len(dv), len(poses), sbits, st0[:, 0].sum(), sbits / max(1, st0[:, 0].sum()), vbits,
                st0[:, 1].sum(), vbits / max(1, st0[:, 1].sum()),
                "   max |d coverage %%| %.4f" % max(dpct) if dpct else ""))
            oc.close()
    out = os.path.join(ROOT, "profiles", "r02_oracle_sensitivity.txt")
    with open(out, "w") as f:
        f.write("\n".join(lines) + "\n")
    print("written", out)


if __name__ == "__main__":
    main()
