"""Drop-in check of the REAL reference file: with `open3d` aliased to b200recon, /root/reference/stitcher.py
imports and RGBDStitcher finds every name it uses (stitcher.py:1-258).  Runs only where /root/reference exists
(the build container); no GPU work is issued."""
import importlib.util
import os
import re
import sys

import pytest

REF = "/root/reference/stitcher.py"


@pytest.mark.skipif(not os.path.exists(REF), reason="/root/reference is not present on this box")
def test_reference_stitcher_binds_to_b200recon(o3d):
    src = open(REF).read()
    used = sorted(set(re.findall(r"\bo3d\.((?:[A-Za-z_][A-Za-z_0-9]*\.)*[A-Za-z_][A-Za-z_0-9]*)", src)))
    assert len(used) > 10
    missing = []
    for name in used:
        obj = o3d
        for part in name.split("."):
            if not hasattr(obj, part):
                missing.append(name)
                break
            obj = getattr(obj, part)
    assert not missing, missing
    # methods called on PointCloud objects in stitch_sequence / preprocess_point_cloud
    for m in ("voxel_down_sample", "estimate_normals", "remove_statistical_outlier", "transform", "has_colors",
              "paint_uniform_color", "__iadd__"):
        assert hasattr(o3d.geometry.PointCloud, m), m
    saved = sys.modules.get("open3d")
    sys.modules["open3d"] = o3d
    try:
        spec = importlib.util.spec_from_file_location("ref_stitcher", REF)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        st = mod.RGBDStitcher(o3d.camera.PinholeCameraIntrinsic(640, 480, 615.67, 615.96, 326.06, 240.56))
        assert st.voxel_size == 0.02 and st.distance_threshold == 0.05
    finally:
        if saved is None:
            sys.modules.pop("open3d", None)
        else:
            sys.modules["open3d"] = saved
