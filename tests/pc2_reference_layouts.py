"""The PointCloud2 messages of the reference's own deserialiser tests (pointmatcher_ros/test/RosPointCloud2DeserializerTest.cpp:195-430):
each test builds a DataPoints (features = Matrix::Random in [-1, 1], pad = 1; optional 'normals', 'color', scalar descriptors), turns
it into a message with pointMatcherCloudToRosMsg<float> (pointmatcher_ros/src/serialization.cpp:11-270) and expects
deserialize(msg) == cloud.  This module restates the SERIALISER's record layout (field order, offsets, point_step) so the same
round trip can be run against the device deserialiser and the oracle; sizes are the reference tests' own."""
import numpy as np

# name, nbPoints, normals, color channels (0 / 3 rgb as FLOAT32 / 4 rgba as UINT32), scalar descriptors, (width, height)
CASES = [
    ("PointCloud3dPoints", 1123455, False, 0, 0, None),                    # :195-219
    ("PointCloud3dPointsNormals", 1867565, True, 0, 0, None),              # :221-247
    ("PointCloud3dPointsColor", 15, False, 3, 0, None),                    # :249-278
    ("PointCloud3dPointsColorWithTransparency", 15, False, 4, 0, None),    # :280-309
    ("PointCloud3dPointsScalar", 10000, False, 0, 1, None),                # :311-337
    ("PointCloud3dPointsNormalsScalars", 10000, True, 0, 5, None),         # :367-397
    ("OrganizedPointCloud3d", 100, False, 0, 0, (25, 4)),                  # :399-430 (width 25 x height 4)
]


def build(case, seed):
    """-> (message bytes uint8[n * point_step], fields {name: offset}, point_step, (width, height), features n x 4, normals n x 3 | None)"""
    name, n, has_n, color, scalars, wh = case
    rng = np.random.default_rng(seed)
    feat = np.ones((n, 4), dtype=np.float32)
    feat[:, :3] = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
    nrm = rng.uniform(-1, 1, (n, 3)).astype(np.float32) if has_n else None
    fields, off, cols = {}, 0, []
    for i, f in enumerate(("x", "y", "z")):                 # serialization.cpp:42-54: features but the padding row
        fields[f] = off
        cols.append(feat[:, i].view(np.uint32))
        off += 4
    if has_n:                                               # :77-99
        for i, f in enumerate(("normal_x", "normal_y", "normal_z")):
            fields[f] = off
            cols.append(nrm[:, i].view(np.uint32))
            off += 4
    if color:                                               # :100-111: ONE 4-byte field, 'rgb' (FLOAT32 datatype) or 'rgba' (UINT32)
        fields["rgba" if color == 4 else "rgb"] = off
        cols.append(rng.integers(0, 2**32, n, dtype=np.uint32))
        off += 4
    for s in range(scalars):                                # :112-120
        fields[f"scalar{s}"] = off
        cols.append(rng.uniform(-1, 1, n).astype(np.float32).view(np.uint32))
        off += 4
    data = np.stack(cols, axis=1).astype(np.uint32).view(np.uint8).reshape(-1)
    return data, fields, off, (wh if wh else (n, 1)), feat, nrm
