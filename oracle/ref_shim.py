"""TEST INFRASTRUCTURE ONLY -- import shim that loads the *unmodified* reference
(ShMonem/animSnapBases, mounted read-only at /root/reference) in this container.

Only `oracle/make_golden.py` and the `-m "not gpu"` pinning tests (which skip when
/root/reference is absent, i.e. on the GPU box) may import this module.  Nothing in
`animsnapbases_b200/` imports anything under `oracle/`.

What the shim does:
  * registers stub modules for GUI/IO packages that are absent here (h5py, igl,
    trimesh, polyscope, matplotlib, mayavi, traits, traitsui, tvtk, pyface).  Import
    time names only -- no graphics API is exercised.
  * loads `utils/utils.py` from source with its last line (the import-time demo call,
    utils/utils.py:938) dropped, and registers it as `utils.utils`.
  * `np.asfarray` shim for utils/process.py:195 (removed in numpy 2).
No reference source is copied into this repository: the files are read from
/root/reference at run time.
"""
from __future__ import annotations

import os
import sys
import types
from unittest import mock

import numpy as np

REFERENCE_ROOT = os.environ.get("ASB_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "snapbases", "posComponents.py"))


class _Stub(types.ModuleType):
    """Module whose every non-dunder attribute is a cached MagicMock."""

    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        m = mock.MagicMock(name=f"{self.__name__}.{name}")
        setattr(self, name, m)
        return m


_STUBS = [
    "h5py", "igl", "igl.copyleft", "igl.copyleft.tetgen", "trimesh", "polyscope", "polyscope.imgui", "matplotlib", "matplotlib.pyplot",
    "matplotlib.cm", "matplotlib.colors", "matplotlib.ticker",
    "mayavi", "mayavi.mlab", "mayavi.tools", "mayavi.tools.mlab_scene_model", "mayavi.core",
    "mayavi.core.ui", "mayavi.core.ui.mayavi_scene", "traits", "traits.api", "traitsui",
    "traitsui.api", "tvtk", "tvtk.api", "tvtk.pyface", "tvtk.pyface.scene_editor", "tvtk.common",
    "pyface", "pyface.timer", "pyface.timer.api", "potpourri3d", "meshio", "tetgen", "pygame",
]

_loaded = False


def load_reference():
    """Make `snapbases.*`, `utils.*`, `config.*` of the reference importable. Idempotent."""
    global _loaded
    if _loaded:
        return
    if not reference_available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    for name in _STUBS:
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = _Stub(name)
    # parent/child wiring so `import a.b` then `a.b` works
    for name in _STUBS:
        if "." in name:
            parent, child = name.rsplit(".", 1)
            if isinstance(sys.modules.get(parent), _Stub):
                setattr(sys.modules[parent], child, sys.modules[name])
    traits_api = sys.modules["traits.api"]
    if isinstance(traits_api, _Stub):
        class HasTraits:  # utils/process.py:293 subclasses it
            pass

        def on_trait_change(*_a, **_k):  # utils/process.py:338 uses it as a decorator factory
            def deco(fn):
                return fn
            return deco

        traits_api.HasTraits = HasTraits
        traits_api.on_trait_change = on_trait_change

    if not hasattr(np, "asfarray"):
        np.asfarray = lambda a, dtype=float: np.asarray(a, dtype=dtype)  # utils/process.py:195

    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)

    import importlib
    utils_pkg = importlib.import_module("utils")  # namespace package
    src_path = os.path.join(REFERENCE_ROOT, "utils", "utils.py")
    with open(src_path, "r") as fh:
        lines = fh.read().rstrip("\n").split("\n")
    # drop the trailing import-time demo call (utils/utils.py:938)
    while lines and (lines[-1].strip() == "" or lines[-1].startswith("run_partitioning_example")):
        lines.pop()
    mod = types.ModuleType("utils.utils")
    mod.__file__ = src_path
    mod.__package__ = "utils"
    sys.modules["utils.utils"] = mod
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        exec(compile("\n".join(lines) + "\n", src_path, "exec"), mod.__dict__)
    setattr(utils_pkg, "utils", mod)
    _loaded = True


def load_pd_module(stem):
    """A module of the reference's consumer package projective_dynamics/ (SURVEY 8(f) n2 / n4), loaded from source under a
    private name.  Those modules say `from utils import ...` and mean projective_dynamics/utils.py, which would shadow the
    root `utils` package registered above: a small stand-in is visible as `utils` only while the module executes."""
    load_reference()
    import importlib.util
    name = "asb_ref_pd_" + stem.lower()
    if name in sys.modules:
        return sys.modules[name]
    saved = sys.modules.get("utils")
    fake = types.ModuleType("utils")
    fake.check_dir_exists = lambda p: os.makedirs(p, exist_ok=True)
    fake.delete_matching_column = mock.MagicMock(name="utils.delete_matching_column")
    sys.modules["utils"] = fake
    try:
        spec = importlib.util.spec_from_file_location(name, os.path.join(REFERENCE_ROOT, "projective_dynamics", stem + ".py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod                      # dataclasses look their module up while the class body executes
        try:
            spec.loader.exec_module(mod)
        except BaseException:
            del sys.modules[name]
            raise
    finally:
        if saved is not None:
            sys.modules["utils"] = saved
        else:
            del sys.modules["utils"]
    return mod


def load_pd_simulators():
    return load_pd_module("Simulators")


class scratch_cwd:
    """`log_time("")` writes function_timings.txt into the CWD (utils/utils.py:232): run the
    reference from a scratch directory."""

    def __init__(self, path="/tmp/asb_oracle_scratch"):
        self.path = path

    def __enter__(self):
        os.makedirs(self.path, exist_ok=True)
        self._old = os.getcwd()
        os.chdir(self.path)
        return self.path

    def __exit__(self, *exc):
        os.chdir(self._old)
        return False
