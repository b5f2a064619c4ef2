"""ORACLE — test / baseline infrastructure only.  Never imported by the product path (vision-transformer_b200/).

Loads the REAL reference (SiddhantSKarki/Vision-Transformer) Python path from the git-ignored staging directory
`oracle/_ref/py/src/` that `oracle/Makefile` fills from the reference tree (`__graft_entry__.build()` runs it; the
directory travels to the GPU box with the snapshot, /root/reference itself does not):

  * `modules()`  -> the reference package src/VisionTransformer/{vit,transformer,config}.py, imported under the
                    alias `ref_VisionTransformer` so that it can live beside the drop-in package of the same name.
                    On a CUDA box this IS the parity oracle and the PyTorch-eager competitor.
  * `train_module(pkg_dir)` -> the reference's src/train.py, UNMODIFIED, executed as a module with `pkg_dir` first
                    on sys.path, i.e. `from VisionTransformer import vit, config` (train.py:21) binds whichever
                    package lives there (the drop-in, or the reference itself for harness checks).

Only tests/, __graft_entry__.smoke() and bench.py's reference / cpu_baseline legs call this.
"""
import importlib.util
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
STAGED = os.path.join(HERE, "_ref", "py", "src")
ALIAS = "ref_VisionTransformer"


def staged_dir():
    """Directory holding the staged reference sources, or None.  The staging copy is preferred; the read-only
    reference tree is used directly only in the build container (it does not exist on the GPU box)."""
    for d in (STAGED, "/root/reference/src"):
        if os.path.exists(os.path.join(d, "VisionTransformer", "vit.py")):
            return d
    return None


def available():
    return staged_dir() is not None


def modules():
    """(vit, transformer, config) modules of the real reference, under the package alias `ref_VisionTransformer`."""
    if ALIAS not in sys.modules:
        src = staged_dir()
        if src is None:
            raise RuntimeError("reference sources are not staged: run `make -C oracle` where /root/reference exists")
        pkg_dir = os.path.join(src, "VisionTransformer")
        spec = importlib.util.spec_from_file_location(ALIAS, os.path.join(pkg_dir, "__init__.py"),
                                                      submodule_search_locations=[pkg_dir])
        pkg = importlib.util.module_from_spec(spec)
        sys.modules[ALIAS] = pkg
        spec.loader.exec_module(pkg)
    vit = importlib.import_module(ALIAS + ".vit")            # its `from . import transformer, config` is relative
    return vit, importlib.import_module(ALIAS + ".transformer"), importlib.import_module(ALIAS + ".config")


def build_model(cfg_kwargs, device="cpu", seed=0):
    """A reference VisionTransformer built the way src/train.py:164-175,65 does, default init under `seed`."""
    import torch
    vit, _, config = modules()
    torch.manual_seed(seed)
    cfg = config.ViTConfig(device=device, **cfg_kwargs)
    return vit.VisionTransformer(cfg).to(device)


def _stub_if_missing(name, attrs=()):
    try:
        importlib.import_module(name)
        return False
    except Exception:
        parts = name.split(".")
        for i in range(1, len(parts) + 1):
            n = ".".join(parts[:i])
            if n not in sys.modules:
                m = types.ModuleType(n)
                m.__path__ = []
                sys.modules[n] = m
                if i > 1:
                    setattr(sys.modules[".".join(parts[:i - 1])], parts[i - 1], m)
        for a in attrs:
            setattr(sys.modules[name], a, None)
        return True


def train_module(pkg_dir, name="ref_train"):
    """Execute the reference's src/train.py unchanged as module `name`, with `pkg_dir` (the directory that CONTAINS
    the package `VisionTransformer`) first on sys.path.  Libraries train.py imports but never uses on the training
    path (matplotlib: train.py:7 — absent from this image) are stubbed in sys.modules only when not importable."""
    src = staged_dir()
    if src is None:
        raise RuntimeError("reference sources are not staged: run `make -C oracle` where /root/reference exists")
    stubbed = [m for m in ("matplotlib.pyplot",) if _stub_if_missing(m)]
    displaced = {}
    for k in [k for k in sys.modules if k == "VisionTransformer" or k.startswith("VisionTransformer.")]:
        if not os.path.abspath(getattr(sys.modules[k], "__file__", "") or "").startswith(os.path.abspath(pkg_dir)):
            displaced[k] = sys.modules.pop(k)      # a different package of that name was imported before: rebind
    sys.path.insert(0, pkg_dir)
    try:
        spec = importlib.util.spec_from_file_location(name, os.path.join(src, "train.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        sys.path.remove(pkg_dir)
    mod.__stubbed__ = stubbed
    mod.__displaced__ = displaced
    return mod


def restore(mod):
    """Put back the `VisionTransformer*` modules that `train_module` displaced from sys.modules."""
    if not mod.__displaced__:
        return
    for k in [k for k in sys.modules if k == "VisionTransformer" or k.startswith("VisionTransformer.")]:
        del sys.modules[k]
    sys.modules.update(mod.__displaced__)
    mod.__displaced__ = {}
