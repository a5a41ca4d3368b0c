"""`o3d.visualization` -- unavailable on this box (no GL/EGL/Vulkan, no software renderer).

The reference only uses it to look at results (stitcher.py:168-200,289; main.py:254).  The calls exist
so that `import b200recon as o3d` keeps `stitcher.py` importable; they do nothing but say so.
"""
import sys


def draw_geometries(geometries, *args, **kwargs):
    sys.stderr.write("[b200recon] visualization.draw_geometries: unavailable on this box (headless); "
                     "%d geometries not shown\n" % len(list(geometries)))


class Visualizer:
    def create_window(self, *a, **k):
        raise RuntimeError("visualization is unavailable on this box (headless)")
