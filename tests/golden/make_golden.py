#!/usr/bin/env python3
"""Generate the committed golden fixtures from the reference's own test data (run in the build container only).

Reads (read-only) from /root/reference/libpointmatcher:
  examples/data/car_cloud400.csv, car_cloud401.csv      -> fixtures of utest/utest.cpp:74-89 (ref3D / data3D)
  examples/data/cloud.00000.vtk, cloud.00001.vtk         -> fixtures of utest/ui/icp/GeneralTests.cpp:70-150,190-210
  examples/data/icp_data/*.ref_trans (on-path ones)      -> golden transforms compared at 5 % (GeneralTests.cpp:136-148)
  utest/utest.cpp:85-89 validT3d                         -> hand-recorded transform, tolerance utest.h:83-84
Writes tests/golden/*.npz / golden_transforms.json.  /root/reference does not exist on the GPU box, so the tests
only ever read the files written here.  These are DATA fixtures, not reference sources.
"""
import json
import os
import numpy as np

REF = "/root/reference/libpointmatcher"
OUT = os.path.dirname(os.path.abspath(__file__))


def load_csv(path):
    with open(path) as f:
        first = f.readline()
    has_header = any(c.isalpha() for c in first.replace("e", "").replace("E", ""))
    if has_header:
        names = [s.strip() for s in first.strip().split(",")]
        data = np.loadtxt(path, delimiter=",", skiprows=1, dtype=np.float64)
    else:
        data = np.loadtxt(path, dtype=np.float64)
        names = ["x", "y", "z"][: data.shape[1]]
    return names, data


def load_vtk_points(path):
    with open(path) as f:
        lines = f.readlines()
    for i, l in enumerate(lines):
        if l.startswith("POINTS"):
            n = int(l.split()[1])
            vals = []
            j = i + 1
            while len(vals) < 3 * n:
                vals.extend(lines[j].split())
                j += 1
            return np.array(vals[: 3 * n], dtype=np.float64).reshape(n, 3)
    raise RuntimeError("no POINTS in " + path)


def main():
    names, d400 = load_csv(f"{REF}/examples/data/car_cloud400.csv")
    assert names == ["x", "y", "z", "nx", "ny", "nz"], names
    _, d401 = load_csv(f"{REF}/examples/data/car_cloud401.csv")
    np.savez_compressed(f"{OUT}/car_clouds.npz",
                        ref_xyz=d400[:, :3].astype(np.float32), ref_normals=d400[:, 3:6].astype(np.float32),
                        data_xyz=d401[:, :3].astype(np.float32))
    c0 = load_vtk_points(f"{REF}/examples/data/cloud.00000.vtk")
    c1 = load_vtk_points(f"{REF}/examples/data/cloud.00001.vtk")
    np.savez_compressed(f"{OUT}/vtk_clouds.npz", cloud0=c0.astype(np.float32), cloud1=c1.astype(np.float32))

    gold = {
        # utest/utest.cpp:85-89
        "validT3d": [[0.982304, 0.166685, -0.0854066, 0.0446816],
                     [-0.150189, 0.973488, 0.172524, 0.191998],
                     [0.111899, -0.156644, 0.981296, -0.0356313],
                     [0, 0, 0, 1]],
    }
    for name in ["defaultPointToPlaneMinDistDataPointsFilter", "force4DOFForPointToPlaneMinimizer"]:
        gold[name] = np.loadtxt(f"{REF}/examples/data/icp_data/{name}.ref_trans").tolist()
    with open(f"{OUT}/golden_transforms.json", "w") as f:
        json.dump(gold, f, indent=1)
    # ICP chain configuration files (data fixtures): the YAML the registrable-module boundary has to load
    ydir = os.path.join(OUT, "yaml")
    os.makedirs(ydir, exist_ok=True)
    import shutil
    for rel in ["libpointmatcher/examples/data/default.yaml", "libpointmatcher/examples/data/default-identity.yaml",
                "libpointmatcher/examples/data/icp_point_to_plane.yaml",
                "libpointmatcher/examples/data/icp_data/defaultPointToPlaneMinDistDataPointsFilter.yaml",
                "libpointmatcher/examples/data/icp_data/force4DOFForPointToPlaneMinimizer.yaml",
                "libpointmatcher/examples/data/unit_tests/badIcpConfig_InvalidModuleType.yaml",
                "libpointmatcher/examples/data/unit_tests/badIcpConfig_InvalidParameter.yaml",
                "open3d_slam_rsl/ros/open3d_slam_ros/param/icp.yaml"]:
        dst = os.path.basename(rel) if "open3d_slam" not in rel else "open3d_slam_icp.yaml"
        shutil.copyfile(os.path.join("/root/reference", rel), os.path.join(ydir, dst))
    print("car_cloud400", d400.shape, "car_cloud401", d401.shape, "cloud0", c0.shape, "cloud1", c1.shape)


if __name__ == "__main__":
    main()
