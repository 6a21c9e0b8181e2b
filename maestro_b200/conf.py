"""Dataset-shape specs for the MAESTRO pretraining path.

The reference describes tiles with hydra-zen dataclasses (``maestro/conf/dataset/*.py``,
``maestro/conf/datasets.py:13-38``) that cannot be imported without hydra-zen and the
raster I/O stack.  The pretraining path only reads a handful of shape fields from them
(``ssl/mim.py:49-79,143-157``, ``ssl/mae.py:55-131``, ``train/model.py:38-50``), so the
product keeps its own plain dataclasses and converts a reference ``DatasetsConfig`` by
duck typing (:func:`spec_from_reference`).  The four shipped dataset shapes are restated
here with the file:line they come from; ``tests/test_oracle_vs_reference.py::test_builtin_specs_equal_reference_configs`` checks them field by field
against the reference configs when /root/reference is present.
"""

from __future__ import annotations

from dataclasses import dataclass, field
from functools import reduce
from math import gcd


@dataclass
class ModalitySpec:
    name: str
    image_size: int                 # px after resize (InputConfig.image_size)
    patch_size: int                 # PatchSizeConfig.mae
    band_groups: list[int]          # channels per band-group G (RasterConfig.bands)
    num_dates: int = 1
    norm_bands: list[int] | None = None   # loss normalisation groups (defaults to band_groups)
    rescale_elev: bool = False
    name_group: str | None = None
    name_embed: str | None = None
    input_size: int | None = None   # px as delivered by the data loader (defaults to image_size)

    @property
    def channels(self) -> int:
        return sum(self.band_groups)

    @property
    def grid(self) -> int:
        return self.image_size // self.patch_size

    @property
    def group(self) -> str:
        return self.name_group if self.name_group is not None else self.name

    @property
    def embed(self) -> str:
        return self.name_embed if self.name_embed else self.name

    @property
    def loss_bands(self) -> list[int]:
        return list(self.norm_bands) if self.norm_bands is not None else list(self.band_groups)

    @property
    def tokens(self) -> int:
        """Tokens of this modality per tile: G * D * g^2."""
        return len(self.band_groups) * self.num_dates * self.grid ** 2

    @property
    def raw_size(self) -> int:
        return self.input_size if self.input_size is not None else self.image_size


@dataclass
class TargetSpec:
    """What the heads read of a reference ``TargetConfig`` / ``TargetRasterConfig`` (conf/dataset/utils.py:49-75;
    ssl/mim.py:169-197).  ``image_size`` (raster targets only) = round(crop_meters / resolution_meters)."""

    name: str
    type_target: str                # "segment" | "classif" | "multilabel_classif"
    num_classes: int
    missing_val: int = -1
    image_size: int | None = None   # px of a raster target; None for per-sample targets

    @property
    def is_raster(self) -> bool:
        return self.image_size is not None


@dataclass
class DatasetSpec:
    name: str
    inputs: dict[str, ModalitySpec]
    grid_pos_enc: int | None = None
    log_inputs: list[str] = field(default_factory=list)
    targets: dict[str, TargetSpec] = field(default_factory=dict)
    ref_input: str | None = None    # modality whose token grid carries raster targets (ssl/mim.py:171-197, 350-372)

    @property
    def groups(self) -> list[tuple[str, str]]:
        """(modality, group) pairs in input order -- conf/dataset/utils.py:111-117."""
        return [(n, m.group) for n, m in self.inputs.items()]

    @property
    def group_names(self) -> list[str]:
        out: list[str] = []
        for _, g in self.groups:
            if g not in out:
                out.append(g)
        return out

    def group_members(self, group: str) -> list[str]:
        return [n for n, g in self.groups if g == group]

    def group_tokens(self, group: str) -> int:
        return sum(self.inputs[n].tokens for n in self.group_members(group))

    @property
    def pos_grid(self) -> int:
        """Side of the sincos table: grid_pos_enc or lcm of the grids (ssl/mim.py:82-88)."""
        if self.grid_pos_enc is not None:
            return self.grid_pos_enc
        return reduce(lambda a, b: a * b // gcd(a, b), [m.grid for m in self.inputs.values()])


def _bands(b) -> list[int]:
    return [b] if isinstance(b, int) else [len(x) for x in b]


def spec_from_reference(datasets) -> DatasetSpec:
    """Convert a reference ``DatasetsConfig`` (or anything with ``.dataset.inputs``)."""
    if isinstance(datasets, DatasetSpec):
        return datasets
    ds = datasets.dataset if hasattr(datasets, "dataset") else datasets
    inputs = {}
    for name, mod in ds.inputs.items():
        inputs[name] = ModalitySpec(
            name=name,
            image_size=int(mod.image_size),
            patch_size=int(mod.patch_size.mae),
            band_groups=_bands(mod.bands),
            num_dates=int(mod.num_dates),
            norm_bands=list(mod.norm_bands) if mod.norm_bands is not None else None,
            rescale_elev=bool(mod.rescale_elev),
            name_group=mod.name_group,
            name_embed=mod.name_embed,
        )
    targets = {}
    for name, t in (getattr(ds, "targets", None) or {}).items():
        res = getattr(t, "resolution_meters", None)
        raster = hasattr(t, "bands") and res is not None           # isinstance(target, RasterConfig), ssl/mim.py:171
        targets[name] = TargetSpec(name, str(t.type_target), int(t.num_classes), int(t.missing_val),
                                   round(ds.crop_meters / res) if raster else None)
    return DatasetSpec(
        name=getattr(datasets, "name_dataset", getattr(ds, "name", "custom")),
        inputs=inputs,
        grid_pos_enc=ds.grid_pos_enc,
        log_inputs=list(getattr(ds, "log_inputs", [])),
        targets=targets,
        ref_input=getattr(ds, "ref_input", None),
    )


def flair_spec() -> DatasetSpec:
    """FLAIR-HUB tile -- conf/dataset/flair.py:33-106 (spot is not in filter_inputs)."""
    return DatasetSpec(
        name="flair",
        grid_pos_enc=160,
        log_inputs=["aerial"],
        targets={"cosia": TargetSpec("cosia", "segment", 15, -1, 512)},       # flair.py:42-44, 113-118 (102.4 m / 0.2 m)
        ref_input="aerial",
        inputs={
            "aerial": ModalitySpec("aerial", 512, 16, [4], norm_bands=[1, 3]),
            "dem": ModalitySpec("dem", 512, 32, [2], rescale_elev=True),
            "s2": ModalitySpec("s2", 10, 2, [10], num_dates=16, norm_bands=[4, 4, 2]),
            "s1_asc": ModalitySpec("s1_asc", 10, 2, [2], num_dates=4, norm_bands=[1, 1],
                                   name_group="s1"),
            "s1_des": ModalitySpec("s1_des", 10, 2, [2], num_dates=4, norm_bands=[1, 1],
                                   name_group="s1"),
        },
    )


def treesatai_ts_spec() -> DatasetSpec:
    """TreeSatAI-TS tile -- conf/dataset/treesatai_ts.py:19-76."""
    return DatasetSpec(
        name="treesatai_ts",
        grid_pos_enc=96,
        log_inputs=["aerial"],
        targets={"treesat_mlc_thresh": TargetSpec("treesat_mlc_thresh", "multilabel_classif", 15, -1)},   # :29-31, 88-92
        inputs={
            "aerial": ModalitySpec("aerial", 300, 20, [4], norm_bands=[1, 3]),
            "s2": ModalitySpec("s2", 6, 2, [10], num_dates=16, norm_bands=[4, 4, 2]),
            "s1_asc": ModalitySpec("s1_asc", 6, 2, [2], num_dates=4, norm_bands=[1, 1],
                                   name_group="s1"),
            "s1_des": ModalitySpec("s1_des", 6, 2, [2], num_dates=4, norm_bands=[1, 1],
                                   name_group="s1"),
        },
    )


def s2_naip_spec() -> DatasetSpec:
    """S2-NAIP-urban tile -- conf/dataset/s2_naip.py:20-84.

    aerial/spot arrive at 96 px (120 m / 1.25 m) and are nearest-resized to 384 / 128
    by ``resize_and_rescale`` (ssl/mim.py:425-437).
    """
    return DatasetSpec(
        name="s2_naip",
        grid_pos_enc=192,
        log_inputs=["aerial", "spot"],
        inputs={
            "aerial": ModalitySpec("aerial", 384, 16, [4], norm_bands=[1, 3], input_size=96),
            "spot": ModalitySpec("spot", 128, 16, [3], input_size=96),
            "s2": ModalitySpec("s2", 12, 2, [10], num_dates=16, norm_bands=[4, 4, 2]),
            "s1": ModalitySpec("s1", 12, 2, [2], num_dates=4, norm_bands=[1, 1]),
        },
    )


def pastis_hd_spec() -> DatasetSpec:
    """PASTIS-HD tile -- conf/dataset/pastis_hd.py:22-75."""
    return DatasetSpec(
        name="pastis_hd",
        grid_pos_enc=256,
        log_inputs=["spot"],
        targets={"pastis_seg": TargetSpec("pastis_seg", "segment", 19, 19, 16)},   # pastis_hd.py:32-34 (160 m / 10 m)
        ref_input="s2",
        inputs={
            "spot": ModalitySpec("spot", 160, 16, [3]),
            "s2": ModalitySpec("s2", 16, 2, [10], num_dates=16, norm_bands=[4, 4, 2]),
            "s1_asc": ModalitySpec("s1_asc", 16, 2, [2], num_dates=4, norm_bands=[1, 1],
                                   name_group="s1"),
            "s1_des": ModalitySpec("s1_des", 16, 2, [2], num_dates=4, norm_bands=[1, 1],
                                   name_group="s1"),
        },
    )


DATASET_SPECS = {
    "flair": flair_spec,
    "treesatai_ts": treesatai_ts_spec,
    "s2_naip": s2_naip_spec,
    "pastis_hd": pastis_hd_spec,
}


@dataclass
class MaskSpec:
    """conf/mask.py:9-15."""

    mask_ratio: float = 0.75
    mask_scale: float = 0.0
    mask_mod: float | None = 0.25
    mask_bands: float | None = None
    mask_dates: float | None = 0.25
    mask_loc: float | None = 0.25


MODEL_SIZES = {
    # ssl/mae.py:309-378
    "tiny": dict(embed_dim=192, depth=12, heads=3, dim_head=64, mlp_ratio=2, decoder_dim=512,
                 decoder_depth=1, decoder_heads=16, decoder_dim_head=32, decoder_mlp_ratio=4),
    "small": dict(embed_dim=384, depth=12, heads=6, dim_head=64, mlp_ratio=2, decoder_dim=512,
                  decoder_depth=2, decoder_heads=16, decoder_dim_head=32, decoder_mlp_ratio=4),
    "medium": dict(embed_dim=768, depth=12, heads=12, dim_head=64, mlp_ratio=4, decoder_dim=512,
                   decoder_depth=3, decoder_heads=16, decoder_dim_head=32, decoder_mlp_ratio=4),
    "large": dict(embed_dim=1024, depth=24, heads=16, dim_head=64, mlp_ratio=4, decoder_dim=512,
                  decoder_depth=4, decoder_heads=16, decoder_dim_head=32, decoder_mlp_ratio=4),
}
