"""Turn the reference's bundled DATA files into compact fixtures under tests/golden/.

Run once in the build container (needs /root/reference, which does not exist on the GPU
box).  Only data is converted (point coordinates, pose tables); no reference source is copied.

  python tools/make_fixtures.py
"""
import os
import sys

import numpy as np

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def read_pcd(path):
    """PCD v0.7 binary, FIELDS x y z float32."""
    with open(path, "rb") as f:
        raw = f.read()
    pos = 0
    header = {}
    while True:
        end = raw.index(b"\n", pos)
        line = raw[pos:end].decode("ascii", "replace").strip()
        pos = end + 1
        if line.startswith("#") or not line:
            continue
        key, _, val = line.partition(" ")
        header[key] = val
        if key == "DATA":
            break
    assert header["DATA"] == "binary" and header["FIELDS"] == "x y z" and header["SIZE"] == "4 4 4"
    n = int(header["POINTS"])
    pts = np.frombuffer(raw, dtype="<f4", count=3 * n, offset=pos).reshape(n, 3).copy()
    vp = np.array([float(t) for t in header["VIEWPOINT"].split()], dtype=np.float64)
    return pts, vp


def read_poses(path, cols=7):
    rows = []
    with open(path) as f:
        for line in f:
            t = line.split()
            if len(t) != cols:
                continue  # the dsscpp file ends in a truncated row
            rows.append([float(x) for x in t])
    return np.array(rows, dtype=np.float64)


def main():
    os.makedirs(OUT, exist_ok=True)
    pts, vp = read_pcd(REF + "/cscpp/pcd/etihad_nowheels_nointernal_scaled_newdensed.pcd")
    print("etihad", pts.shape, "viewpoint", vp, "bbox", pts.min(0), pts.max(0))
    np.savez_compressed(os.path.join(OUT, "etihad_model.npz"), xyz=pts, viewpoint=vp)

    t = REF + "/cscpp/txt/"
    uav = read_poses(t + "SearchSpaceUAV_2.5m_1to4_etihad_nowheels_nointernal_scaled_newdensed.pcd.txt")
    cam0 = read_poses(t + "SearchSpaceCam_2.5m_1to4_etihad_nowheels_nointernal_scaled_newdensed.pcd_0.txt")
    cam1 = read_poses(t + "SearchSpaceCam_2.5m_1to4_etihad_nowheels_nointernal_scaled_newdensed.pcd_1.txt")
    dss = read_poses(t + "search_space_robot_etihad_nowheels_nointernal_scaled_newdensed.pcd_dsscpp.txt", cols=4)
    print("2.5m search space", uav.shape, cam0.shape, cam1.shape, "dsscpp", dss.shape)
    np.savez_compressed(os.path.join(OUT, "etihad_searchspace_2p5m.npz"), uav=uav, cam0=cam0, cam1=cam1, dsscpp=dss)

    r = REF + "/sspp/resources/"
    uav = read_poses(r + "SearchSpaceUAV_1.5m_1to4_etihad_nowheels_nointernal_newdensed.txt")
    cam0 = read_poses(r + "SearchSpaceCam_1.5m_1to4_etihad_nowheels_nointernal_newdensed_0.txt")
    cam1 = read_poses(r + "SearchSpaceCam_1.5m_1to4_etihad_nowheels_nointernal_newdensed_1.txt")
    print("1.5m search space", uav.shape, cam0.shape, cam1.shape)
    np.savez_compressed(os.path.join(OUT, "etihad_searchspace_1p5m.npz"), uav=uav, cam0=cam0, cam1=cam1)


if __name__ == "__main__":
    sys.exit(main())
