"""TEST INFRASTRUCTURE ONLY -- import shim for the real reference under /root/reference.

This file lets the *unmodified* reference modules (``maestro.ssl.mae.MAE``,
``maestro.train.model.SSLModule.compute_loss_rec``) be imported in the build container,
where several third-party packages the reference imports are absent.  Every stub
installed here is arithmetic-free (config plumbing, Lightning base classes, raster I/O);
the only stub that carries arithmetic is ``vit_pytorch.vit`` which is served by the
restatement in ``oracle/vit_restated.py`` (vit-pytorch 1.10.1 is a third-party
dependency pinned in the reference's poetry.lock:6339-6347 and is not vendored).

It is used by ``tests/golden/make_golden.py`` (to generate the committed fixtures) and
by ``tests/test_oracle_vs_reference.py`` (skipped when /root/reference is absent, e.g.
on the GPU box).  Nothing in the product package imports this file.
"""

from __future__ import annotations

import importlib
import importlib.machinery
import os
import sys
import types

def _reference_root() -> str:
    """/root/reference in the build container; on the GPU box the mirror made by ``oracle/make_ref.py``
    (``oracle/_ref/``: git-ignored, shipped with the gpurun snapshot)."""
    env = os.environ.get("MAESTRO_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isdir("/root/reference/maestro"):
        return "/root/reference"
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


REFERENCE_ROOT = _reference_root()


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "maestro"))


class _Anything:
    """Permissive placeholder: callable, subscriptable, attribute-able."""

    def __init__(self, *a, **k) -> None:
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Anything()

    def __getitem__(self, k):
        return _Anything()

    def __iter__(self):
        return iter(())


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        cls = type(name, (_Anything,), {})
        setattr(self, name, cls)
        return cls


class _NNStubModule(_StubModule):
    """Stub module whose attributes are permissive nn.Module subclasses (torchmetrics)."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        from torch import nn

        def _init(self, *a, **k):
            nn.Module.__init__(self)

        cls = type(name, (nn.Module,), {"__init__": _init,
                                        "add_state": lambda self, *a, **k: None,
                                        "update": lambda self, *a, **k: None,
                                        "compute": lambda self, *a, **k: None,
                                        "reset": lambda self, *a, **k: None})
        setattr(self, name, cls)
        return cls


_NN_TOPS = {"torchmetrics"}


def _stub(name: str) -> types.ModuleType:
    if name in sys.modules and not isinstance(sys.modules[name], _StubModule):
        return sys.modules[name]
    mod = (_NNStubModule if name.split(".")[0] in _NN_TOPS else _StubModule)(name)
    mod.__path__ = []  # behave as a package so submodule imports resolve
    mod.__spec__ = importlib.machinery.ModuleSpec(name, loader=None, is_package=True)
    sys.modules[name] = mod
    parent, _, child = name.rpartition(".")
    if parent:
        setattr(_stub(parent), child, mod)
    return mod


class _StubFinder:
    """Resolve any not-yet-seen submodule of a stubbed top-level package to a stub."""

    @staticmethod
    def find_spec(name, path=None, target=None):
        top = name.split(".")[0]
        if isinstance(sys.modules.get(top), _StubModule) and name not in sys.modules:
            return importlib.machinery.ModuleSpec(name, _StubFinder, is_package=True)
        return None

    @staticmethod
    def create_module(spec):
        cls = _NNStubModule if spec.name.split(".")[0] in _NN_TOPS else _StubModule
        mod = cls(spec.name)
        mod.__path__ = []
        return mod

    @staticmethod
    def exec_module(mod):
        return None


def _have(name: str) -> bool:
    try:
        importlib.import_module(name)
        return True
    except Exception:  # noqa: BLE001
        return False


def install() -> None:
    """Install stubs and put the reference on sys.path. Idempotent."""
    import torch
    from torch import nn

    if getattr(install, "_done", False):
        return
    if _StubFinder not in sys.meta_path:
        sys.meta_path.append(_StubFinder)
    if not reference_available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")

    # ---- hydra_zen: only MISSING / store / builds are touched at import time ----
    if not _have("hydra_zen"):
        hz = _stub("hydra_zen")
        hz.MISSING = "???"

        class _Store:
            def __call__(self, *a, **k):
                if a and callable(a[0]) and not k.get("group"):
                    return a[0]
                return _Store()

            def add_to_hydra_store(self, *a, **k):
                return None

        hz.store = _Store()
        hz.builds = lambda *a, **k: _Anything()
        hz.zen = lambda *a, **k: _Anything()

    # ---- Lightning / torchmetrics: base classes only ----
    if not _have("pytorch_lightning"):
        pl = _stub("pytorch_lightning")

        class LightningModule(nn.Module):
            def save_hyperparameters(self, *a, **k):
                return None

            def log(self, *a, **k):
                return None

        pl.LightningModule = LightningModule
        pl.Trainer = type("Trainer", (), {})
        pl.Callback = type("Callback", (), {})
        pl.LightningDataModule = type("LightningDataModule", (), {})
        for sub in ("callbacks", "loggers", "utilities", "loggers.tensorboard",
                    "utilities.types", "utilities.rank_zero"):
            _stub(f"pytorch_lightning.{sub}")
    if not _have("torchmetrics"):
        tm = _stub("torchmetrics")

        class MeanMetric(nn.Module):
            def __init__(self, *a, **k):
                super().__init__()
                self.values = []

            def update(self, v):
                self.values.append(float(v))

        class _Metric(nn.Module):
            def __init__(self, *a, **k):
                super().__init__()

        tm.MeanMetric = MeanMetric
        tm.Metric = _Metric
        tm.MetricCollection = _Metric
        for sub in ("classification", "functional", "functional.classification"):
            _stub(f"torchmetrics.{sub}")

    # ---- raster / geo / plotting I/O: never executed on the hot path ----
    for name in ("rasterio", "rasterio.windows", "rasterio.enums", "h5py", "geopandas",
                 "matplotlib", "matplotlib.pyplot", "matplotlib.colors", "matplotlib.cm",
                 "matplotlib.figure", "seaborn", "clearml", "dacite", "omegaconf",
                 "hydra", "hydra.core", "hydra.core.hydra_config", "hydra.conf",
                 "timm", "terratorch", "torchgeo", "torchvision",
                 "torchvision.utils", "torchvision.transforms",
                 "torchvision.transforms.functional"):
        if not _have(name):
            _stub(name)
    if isinstance(sys.modules.get("rasterio"), _StubModule):
        errs = _stub("rasterio.errors")
        errs.NotGeoreferencedWarning = type("NotGeoreferencedWarning", (Warning,), {})

    # ---- vit_pytorch.vit: restated (third-party, absent) ----
    if not _have("vit_pytorch"):
        here = os.path.dirname(os.path.abspath(__file__))
        if here not in sys.path:
            sys.path.insert(0, here)
        import vit_restated  # noqa: PLC0415

        vp = types.ModuleType("vit_pytorch")
        vp.__path__ = []
        vit = types.ModuleType("vit_pytorch.vit")
        vit.Transformer = vit_restated.Transformer
        vit.Attention = vit_restated.Attention
        vit.FeedForward = vit_restated.FeedForward
        vp.vit = vit
        sys.modules["vit_pytorch"] = vp
        sys.modules["vit_pytorch.vit"] = vit

    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    torch.set_float32_matmul_precision("highest")
    install._done = True


def reference_datasets_config(name_dataset: str):
    """Build the reference's own DatasetsConfig for one of its shipped datasets."""
    install()
    from maestro.conf.dataset.flair import FLAIRConfig
    from maestro.conf.dataset.pastis_hd import PASTISHDConfig
    from maestro.conf.dataset.s2_naip import S2NAIPConfig
    from maestro.conf.dataset.treesatai_ts import TreeSatAITSConfig
    from maestro.conf.datasets import DatasetsConfig

    return DatasetsConfig(
        root_dir="/nonexistent",
        name_dataset=name_dataset,
        treesatai_ts=TreeSatAITSConfig(),
        pastis_hd=PASTISHDConfig(),
        flair=FLAIRConfig(),
        s2_naip=S2NAIPConfig(),
    )
