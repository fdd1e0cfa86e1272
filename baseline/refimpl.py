"""The UNMODIFIED reference, made importable for bench.py's reference arm and side baselines.

pem-gini/GndNet has no setup.py / pyproject.toml, so `pip install --target baseline/_ref /root/reference`
does not apply.  The offline install is a verbatim copy of the reference's Python packages into the
git-ignored `baseline/_ref/gnd_net/` (run by `__graft_entry__.build()` in the build container, where
/root/reference exists); `baseline/_ref/` is NOT gpurun-ignored, so the copy travels to the GPU box.
Nothing of it is committed, and nothing under `gndnet_b200/` imports it.

`load()` imports it as the package `gnd_net` -- the name the reference uses for itself (model.py:8-9) -- with empty
stand-ins for the plotting / ROS imports at the top of utils/utils.py:22-31 that this image does not have
(matplotlib, shapely, ros2_numpy: none of them is on the hot path).
"""
import importlib
import os
import shutil
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference"
DST = os.path.join(ROOT, "baseline", "_ref", "gnd_net")
# what the hot path imports (model.py, modules/*, torchplus/*, utils/*) plus the two shipped configs
_TOP_FILES = ["__init__.py", "model.py"]
_DIRS = ["modules", "utils", "torchplus", "config"]


def install(force=False):
    """Copy the reference's hot-path packages verbatim.  No-op without /root/reference (the GPU box)."""
    if not os.path.isdir(SRC):
        return os.path.isdir(DST) and not os.path.islink(DST)
    if os.path.islink(DST):                    # round 1 left a symlink that does not resolve on the GPU box
        os.unlink(DST)
    if os.path.isdir(DST) and not force:
        return True
    if os.path.isdir(DST):
        _make_writable(DST)
        shutil.rmtree(DST)
    os.makedirs(DST)
    for f in _TOP_FILES:
        shutil.copy2(os.path.join(SRC, f), os.path.join(DST, f))
    for d in _DIRS:
        shutil.copytree(os.path.join(SRC, d), os.path.join(DST, d),
                        ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    _make_writable(DST)                        # /root/reference is read-only; the copy must stay removable
    return True


def _make_writable(top):
    for d, _, files in os.walk(top):
        os.chmod(d, 0o755)
        for f in files:
            os.chmod(os.path.join(d, f), 0o644)


def available():
    return os.path.isfile(os.path.join(DST, "model.py"))


def load():
    """-> (point_cloud_ops, model, utils) modules of the unmodified reference.  Raises ImportError if the copy or
    one of its dependencies (numba) is missing."""
    if not available():
        raise ImportError("baseline/_ref/gnd_net is not installed (run __graft_entry__.build() where /root/reference exists)")
    os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
    parent = os.path.dirname(DST)
    if parent not in sys.path:
        sys.path.insert(0, parent)
    for name in ["matplotlib", "matplotlib.pyplot", "shapely", "shapely.geometry",
                 "gnd_net.ros2_numpy", "gnd_net.ros2_numpy.ros2_numpy"]:
        if name not in sys.modules:
            try:
                if not name.startswith("gnd_net"):
                    importlib.import_module(name)
                    continue
            except ImportError:
                pass
            m = types.ModuleType(name)
            if name == "shapely.geometry":
                m.Polygon = object
                m.Point = object
            sys.modules[name] = m
    ops = importlib.import_module("gnd_net.utils.point_cloud_ops")
    model = importlib.import_module("gnd_net.model")
    utils = importlib.import_module("gnd_net.utils.utils")
    return ops, model, utils
