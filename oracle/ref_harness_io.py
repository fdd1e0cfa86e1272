"""BUILD-CONTAINER ONLY: pins the SURVEY.md 8(f) rows (cloud unpack / transform / NaN filter upstream, wire packing
downstream) of the oracle against the unmodified reference, and writes tests/golden/cloud_io.npz.

    python -m oracle.ref_harness_io [--write]

The reference functions run as they are: utils.utils.cloud_msg_to_numpy on a stand-in message object (fields,
data, point_step -- the only attributes it reads), utils.transform.transformStampedToTransformationMatrix +
transformCloud (numba), the NaN filter expression of ros_node.py:263, utils.ros_utils.array_to_pointcloud2 and
the structured array np2ros_pub_2 hands to ros2_numpy.msgify (ROS message classes stubbed: absent here).
"""
import argparse
import os
import sys
import types
from types import SimpleNamespace as NS

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def import_reference_io():
    from oracle.ref_harness import import_reference
    import_reference()

    class _Msg:                      # sensor_msgs.msg.PointCloud2 / PointField stand-ins: plain attribute bags
        def __init__(self, **kw):
            self.__dict__.update(kw)
    _Msg.FLOAT32 = 7
    for name, attrs in [("sensor_msgs", {}), ("sensor_msgs.msg", {"PointCloud2": _Msg, "PointField": _Msg}),
                        ("std_msgs", {}), ("std_msgs.msg", {"Header": _Msg}),
                        ("visualization_msgs", {}), ("visualization_msgs.msg", {"Marker": _Msg, "MarkerArray": _Msg}),
                        ("geometry_msgs", {}), ("geometry_msgs.msg", {"Point": _Msg}),
                        ("torchvision", {"datasets": None, "transforms": None})]:
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__dict__.update(attrs)
            sys.modules[name] = m
    import importlib
    utils = importlib.import_module("gnd_net.utils.utils")
    transform = importlib.import_module("gnd_net.utils.transform")
    try:
        ros_utils = importlib.import_module("gnd_net.utils.ros_utils")
    except Exception as ex:          # packing is then pinned on the structured-array restatement only
        print("ros_utils import failed:", repr(ex))
        ros_utils = None
    return utils, transform, ros_utils


def make_msg(rng, n, point_step, with_nan=True, with_intensity=True):
    """A PointCloud2-like byte buffer: x, y, z float32 at offsets 0, 4, 8; the rest of the row is other fields."""
    data = rng.integers(0, 256, size=(n, point_step), dtype=np.uint8)
    xyz = (rng.random((n, 3)) * np.array([120.0, 120.0, 8.0]) - np.array([60.0, 60.0, 5.0])).astype(np.float32)
    if with_nan:
        bad = rng.random(n) < 0.03
        col = rng.integers(0, 3, size=n)
        xyz[bad, col[bad]] = np.nan
        xyz[rng.random(n) < 0.002, 0] = np.inf
    data[:, :12] = xyz.view(np.uint8).reshape(n, 12)
    names = ["x", "y", "z"] + (["intensity"] if with_intensity and point_step >= 18 else [])
    fields = [NS(name=k) for k in names]
    return NS(data=data.tobytes(), point_step=point_step, fields=fields, header=NS(frame_id="velo")), data


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--write", action="store_true")
    args = ap.parse_args()
    sys.path.insert(0, ROOT)
    from oracle import cpu_ops
    utils, transform, ros_utils = import_reference_io()
    rng = np.random.default_rng(20260101)
    gold = {}
    cases = [("step16", 8192, 16, True), ("step18_unaligned", 20011, 18, True), ("step32", 120000, 32, True),
             ("step12_identity", 4096, 12, False), ("empty", 0, 16, True), ("all_nan", 64, 16, True)]
    for name, n, step, use_t in cases:
        msg, raw = make_msg(rng, n, step)
        if name == "all_nan":
            raw[:, :4] = np.frombuffer(np.float32(np.nan).tobytes(), dtype=np.uint8)
            msg.data = raw.tobytes()
        ts = NS(transform=NS(translation=NS(x=1.25, y=-0.4, z=1.733),
                             rotation=NS(x=0.0174524, y=-0.0348995, z=0.3826834, w=0.9231444)))
        trafo = transform.transformStampedToTransformationMatrix(ts) if use_t else None
        # ---- the reference, as ros_node.py:242-263 runs it
        if n > 0:
            cloud = utils.cloud_msg_to_numpy(msg, 0, shift_cloud=False)["xyz"]
            cloud = np.array(cloud)                   # frombuffer views are read-only; the node's message buffer is not
            if trafo is not None:
                cloud = transform.transformCloud(cloud, trafo)
            cloud = cloud[~np.isnan(cloud).any(axis=1)]
        else:
            cloud = np.zeros((0, 3), np.float32)
        # ---- the oracle
        mine = cpu_ops.unpack_cloud(raw, step, trafo, True, 3)
        same = mine.shape == cloud.shape and np.array_equal(mine.view(np.uint32), np.ascontiguousarray(cloud).view(np.uint32))
        print("%-18s n=%6d kept=%6d  unpack/transform/nan-filter bit-exact: %s" % (name, n, len(cloud), same))
        assert same, name
        if n <= 21000:
            gold[name + "__msg"] = raw
            gold[name + "__trafo"] = np.zeros(0) if trafo is None else np.asarray(trafo, dtype=np.float64)
            gold[name + "__cloud"] = np.ascontiguousarray(cloud)
    # ---- downstream packing
    n = 6000
    pts = (rng.random((n, 3)) * 100 - 50).astype(np.float32)
    lab = rng.integers(-1, 2, size=n).astype(np.float64)
    arr = np.zeros((n,), dtype=[("x", np.float32), ("y", np.float32), ("z", np.float32), ("intensity", np.float32)])
    t = np.transpose(pts)                              # np2ros_pub_2, ros_utils.py:171-185, up to the msgify call
    arr["x"], arr["y"], arr["z"] = t[0], t[1], t[2]
    arr["intensity"] = lab
    wire16 = np.frombuffer(arr.tobytes(), dtype=np.uint8)
    mine16 = cpu_ops.pack_segmented(pts, lab)
    ok16 = np.array_equal(np.frombuffer(mine16.tobytes(), dtype=np.uint8), wire16)
    print("np2ros_pub_2 16-byte rows bit-exact:", ok16)
    assert ok16
    obs = pts[lab == 1]
    if ros_utils is not None:
        class _A(np.ndarray):            # ndarray.tostring (ros_utils.py:247) is gone from numpy 2: same bytes as tobytes
            def tostring(self):
                return self.tobytes()
        m = ros_utils.array_to_pointcloud2(obs.view(_A), NS(stamp=0, frame_id="map"))
        ok12 = bytes(m._data) == np.ascontiguousarray(obs).tobytes() and m.point_step == 12 and m.width == len(obs)
        print("array_to_pointcloud2 12-byte rows == compacted obstacle rows:", ok12)
        assert ok12
    gold["pack__points"], gold["pack__labels"], gold["pack__wire16"] = pts, lab, wire16
    # ---- evaluation reductions (8f row 3): the reference's own get_GndSeg / remove_outliers, lifted out of the
    # script evaluate_SemanticKITTI.py (module-level argparse / rclpy: not importable) by exec'ing exactly those defs
    import ast
    src = open("/root/reference/evaluate_SemanticKITTI.py").read()
    ns = {"np": np}
    from numba import jit
    ns["jit"] = jit
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name in ("get_GndSeg", "remove_outliers"):
            exec(compile(ast.Module([node], []), "evaluate_SemanticKITTI.py", "exec"), ns)
    n = 40000
    sem = rng.choice(np.array([0, 1, 10, 40, 44, 48, 49, 50, 60, 70, 72, 80, 252], dtype=np.uint32), size=n)
    pred = rng.integers(-1, 2, size=n).astype(np.float64)
    GndSeg = ns["get_GndSeg"](sem, GndClasses=[40, 44, 48, 49, 60, 72])
    p2, g2 = ns["remove_outliers"](pred, GndSeg)
    inter, union = np.logical_and(g2, p2), np.logical_or(g2, p2)
    ref_counts = np.array([np.sum(inter), np.sum(union), p2.sum(), g2.sum()], dtype=np.int64)
    ref_metrics = np.array([np.sum(inter) / np.sum(union), np.sum(inter) / p2.sum(), np.sum(inter) / g2.sum()])
    mine = cpu_ops.ground_counts(pred, sem)
    print("evaluation counts (I, U, P, G):", ref_counts.tolist(), "oracle equal:", np.array_equal(mine, ref_counts))
    assert np.array_equal(mine, ref_counts)
    gold["eval__sem"], gold["eval__pred"], gold["eval__counts"], gold["eval__metrics"] = sem, pred.astype(np.int8), ref_counts, ref_metrics
    # ---- ground-truth rasterisation (8f row 3): the reference's numba lidar_to_img / lidar_to_heightmap as they are
    from gndnet_b200 import synth
    cfgk = synth.make_cfg("kitti")
    gr = np.asarray(cfgk.grid_range)
    with np.load(os.path.join(GOLDEN, "real_000000.npz")) as z:
        real = z["points"].astype(np.float32)[:, :3].copy()       # cells with up to ~1000 points: the overrun is exercised
    hd = synth.hdl64_scan(3)[:, :3].copy()
    hd[:, 2] += np.float32(1.733)
    for name, pts, mp in (("real", real, 100), ("hdl64", hd, 100), ("hdl64_mp5", hd, 5)):
        r_img = utils.lidar_to_img(np.copy(pts), gr, cfgk.voxel_size[0], fill=1)
        r_h = utils.lidar_to_heightmap(np.copy(pts), gr, cfgk.voxel_size[0], max_points=mp)
        o_img = cpu_ops.lidar_to_img(pts, gr, cfgk.voxel_size[0], 1)
        o_h = cpu_ops.lidar_to_heightmap(pts, gr, cfgk.voxel_size[0], mp)
        d = float(np.abs(r_h - o_h).max())
        print("rasterisation %-10s img equal %s, heightmap max|d| %.3g" % (name, np.array_equal(r_img, o_img), d))
        assert np.array_equal(r_img, o_img) and d <= 1e-12
        gold["raster__%s__img" % name], gold["raster__%s__height" % name] = r_img.astype(np.uint8), r_h
    if args.write:
        np.savez_compressed(os.path.join(GOLDEN, "cloud_io.npz"), **gold)
        print("wrote tests/golden/cloud_io.npz", sum(v.nbytes for v in gold.values()) // 1024, "KiB raw")


if __name__ == "__main__":
    main()
