"""Stand-in for the `plyfile` package, which the reference's scene/gaussian_model.py and scene/dataset_readers.py import at
module level and which is not installed in this image.  tests/test_reference_callers.py puts this directory
on sys.path so the UNMODIFIED reference modules import; nothing on the tested path (render, create_from_pcd) touches PLY
files, and any use raises.  (The product's own PLY reader/writer is rain_b200/ply.py.)"""


class _Unavailable:
    def __init__(self, *a, **k):
        raise RuntimeError("plyfile is not installed in this image; use rain_b200.ply")

    @classmethod
    def read(cls, *a, **k):
        raise RuntimeError("plyfile is not installed in this image; use rain_b200.ply")

    @classmethod
    def describe(cls, *a, **k):
        raise RuntimeError("plyfile is not installed in this image; use rain_b200.ply")


class PlyData(_Unavailable):
    pass


class PlyElement(_Unavailable):
    pass
