"""Import the reference's own hot-path modules, unmodified, from /root/reference.

TEST INFRASTRUCTURE (see oracle/__init__.py).  The whole tree is mounted only in the
build container; the GPU box has no /root/reference, but the three hot-path files
travel there as oracle/_ref/ (copied unmodified by oracle/make_ref.py, git-ignored).
Used by tests/golden/make_golden.py to generate the committed golden vectors, by the
`-m "not gpu"` cross-check of the restatement, and by bench.py's CPU arms (front_view /
bev_density only -- the train/predict.py helpers below need the full tree).

The reference modules import ROS / matplotlib at module scope
(modules/lidar/process/extract_rosbag_lidar.py:10-16,
modules/lidar/common/birds_eye_view_generator.py:1-9); none of those is needed by
the numeric functions, so empty stand-in modules are registered first.  For the BEV
the only thing matplotlib does is the final colormap (birds_eye_view_generator.py:
48-53); a stand-in ScalarMappable captures the float array handed to `to_rgba`,
which is the pre-colormap density image the parity contract is written against.
"""
import importlib.util
import os
import sys
import types

def _default_root():
    """The mounted reference tree (build container) or, where that does not exist (the GPU box), the three
    hot-path files that oracle/make_ref.py copied unmodified into oracle/_ref/."""
    env = os.environ.get("SFPE_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isdir("/root/reference/modules/lidar"):
        return "/root/reference"
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


REFERENCE_ROOT = _default_root()
_LIDAR = os.path.join(REFERENCE_ROOT, "modules", "lidar")

_cache = {}


def available():
    """The three hot-path files (front view, BEV, constants) can be loaded."""
    return os.path.isfile(os.path.join(_LIDAR, "process", "extract_rosbag_lidar.py"))


def full_tree_available():
    """train/predict.py is there too (only in the build container, never in oracle/_ref)."""
    return available() and os.path.isfile(os.path.join(_LIDAR, "train", "predict.py"))


class _Capture:
    """Holds the last array passed to the stand-in colormap."""
    last = None


def _install_stubs():
    def mod(name):
        m = sys.modules.get(name)
        if m is None:
            m = types.ModuleType(name)
            sys.modules[name] = m
        return m

    mod("rosbag")
    sm = mod("sensor_msgs")
    sm.point_cloud2 = mod("sensor_msgs.point_cloud2")
    mpl = mod("matplotlib")
    if not hasattr(mpl, "use"):
        mpl.use = lambda *a, **k: None
    for sub in ("cm", "colors", "image", "pyplot"):
        setattr(mpl, sub, mod("matplotlib." + sub))

    class Normalize:
        def __init__(self, vmin=None, vmax=None):
            self.vmin, self.vmax = vmin, vmax

    class ScalarMappable:
        def __init__(self, cmap=None, norm=None):
            self.cmap, self.norm = cmap, norm

        def to_rgba(self, x, bytes=False, norm=True):
            import numpy as np
            _Capture.last = np.array(x, copy=True)
            return np.zeros(x.shape + (4,), dtype=np.uint8)

    if not hasattr(mpl.colors, "Normalize"):
        mpl.colors.Normalize = Normalize
    if not hasattr(mpl.cm, "ScalarMappable"):
        mpl.cm.ScalarMappable = ScalarMappable


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def load():
    """Returns (extract_rosbag_lidar module, birds_eye_view_generator module, globals module)."""
    if "mods" in _cache:
        return _cache["mods"]
    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    _install_stubs()
    proc = os.path.join(_LIDAR, "process")
    # `import globals` (py2-style sibling import) must resolve to process/globals.py
    sys.path.insert(0, proc)
    try:
        sys.modules.pop("globals", None)
        erl = _load("_ref_extract_rosbag_lidar", os.path.join(proc, "extract_rosbag_lidar.py"))
        bev = _load("_ref_birds_eye_view_generator",
                    os.path.join(_LIDAR, "common", "birds_eye_view_generator.py"))
        glb = sys.modules["globals"]
    finally:
        sys.path.remove(proc)
    _cache["mods"] = (erl, bev, glb)
    return _cache["mods"]


def front_view(points):
    """reference generate_lidar_2d_front_view(points, cmap=None) -> dict of float64 (32,1801)."""
    erl, _, _ = load()
    return erl.generate_lidar_2d_front_view(points, cmap=None)


def bev_density(points, res, max_range=120):
    """Pre-colormap float64 density image of reference generate_birds_eye_view."""
    _, bev, _ = load()
    _Capture.last = None
    bev.generate_birds_eye_view(points, 0, res, max_range=max_range, cmap="gray")
    return _Capture.last


def back_project_fn():
    """The reference's back_project_2D_2_3D (train/predict.py:230-293), unmodified.  predict.py as a whole
    needs rosbag, cv2, keras and python-2 print statements, so only that function's own source lines are
    compiled, in a namespace holding exactly the names it uses (numpy, math.sin/cos and the constants of
    process/globals.py)."""
    if "bp" in _cache:
        return _cache["bp"]
    _, _, glb = load()
    path = os.path.join(_LIDAR, "train", "predict.py")
    with open(path) as f:
        lines = f.readlines()
    start = next(i for i, ln in enumerate(lines) if ln.startswith("def back_project_2D_2_3D("))
    end = next(i for i in range(start + 1, len(lines)) if lines[i].startswith(("def ", "class ")))
    import math
    import numpy as np
    ns = {"np": np, "sin": math.sin, "cos": math.cos, "X_MIN": glb.X_MIN, "RES_RAD": glb.RES_RAD,
          "LIDAR_MIN_HEIGHT": glb.LIDAR_MIN_HEIGHT}
    exec(compile("".join(lines[start:end]), path, "exec"), ns)
    _cache["bp"] = ns["back_project_2D_2_3D"]
    return _cache["bp"]


def _predict_functions(names):
    """Compile the named top-level functions of train/predict.py (and the module constants they read) from their
    own source lines -- see back_project_fn for why the module cannot be imported whole."""
    _, _, glb = load()
    path = os.path.join(_LIDAR, "train", "predict.py")
    with open(path) as f:
        lines = f.readlines()
    import math
    import numpy as np
    ns = {"np": np, "sin": math.sin, "cos": math.cos, "atan2": math.atan2, "pi": math.pi, "X_MIN": glb.X_MIN,
          "Y_MIN": glb.Y_MIN, "RES_RAD": glb.RES_RAD, "LIDAR_MIN_HEIGHT": glb.LIDAR_MIN_HEIGHT}
    for ln in lines:      # MIN_PROB, MIN_BBOX_AREA, MIN_HEAT, MAX_BBX_DIST (predict.py:28-31)
        if ln.startswith(("MIN_PROB", "MIN_BBOX_AREA", "MIN_HEAT", "MAX_BBX_DIST")):
            exec(ln, ns)
    for name in names:
        start = next(i for i, ln in enumerate(lines) if ln.startswith("def %s(" % name))
        end = next(i for i in range(start + 1, len(lines)) if lines[i].startswith(("def ", "class ")))
        exec(compile("\n" * start + "".join(lines[start:end]), path, "exec"), ns)
    return ns


def find_bbox_3d_fn():
    """The reference's find_bbox_3d (train/predict.py:94-199) with its helper is_far (:83-92), unmodified."""
    if "fb" not in _cache:
        _cache["fb"] = _predict_functions(["is_far", "find_bbox_3d"])["find_bbox_3d"]
    return _cache["fb"]


def find_obstacle_fn():
    """The reference's find_obstacle (train/predict.py:33-81), unmodified, with scipy.ndimage.label as it imports it
    (:16).  It spells the dtype `np.float`, an alias numpy 2 removed, so the name `np` it sees is numpy plus that one
    alias (= the builtin float, which is what the alias was)."""
    if "fo" in _cache:
        return _cache["fo"]
    import numpy as np
    from scipy.ndimage import label

    class _NumpyWithFloatAlias:
        float = float

        def __getattr__(self, name):
            return getattr(np, name)

    ns = _predict_functions([])
    ns.update(np=_NumpyWithFloatAlias(), label=label)
    path = os.path.join(_LIDAR, "train", "predict.py")
    with open(path) as f:
        lines = f.readlines()
    start = next(i for i, ln in enumerate(lines) if ln.startswith("def find_obstacle("))
    end = next(i for i in range(start + 1, len(lines)) if lines[i].startswith(("def ", "class ")))
    exec(compile("\n" * start + "".join(lines[start:end]), path, "exec"), ns)
    _cache["fo"] = ns["find_obstacle"]
    return _cache["fo"]
