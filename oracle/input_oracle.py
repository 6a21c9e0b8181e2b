"""TEST INFRASTRUCTURE ONLY -- CPU (numpy) restatement of the input pipeline the device kernels of
``maestro_b200/csrc/input.cu`` replace.  Only tests/, ``__graft_entry__.smoke()`` and bench.py's CPU leg may import it.

* ``preprocess_tail`` follows ``GenericDataset.preprocess_raster`` maestro/dataset/dataset.py:215-222.
* ``transform_rasters`` follows ``GenericDataset.transform_rasters`` maestro/dataset/dataset.py:225-257.
  Both are pinned by ``tests/golden/input_pipeline.npz``, generated from the reference's own methods by
  ``tests/golden/make_input_golden.py``.
* ``struct_mask`` restates ``MAE.mask_struct`` maestro/ssl/mae.py:178-226 per sample with the counter-based
  Philox4x32-10 stream the device kernel uses (performance mode: same distribution as the reference's host
  ``torch.rand`` loop, different stream).  ``philox4x32_10`` is pinned by the Random123 known-answer vectors
  (tests/test_input_oracle.py).
"""

from __future__ import annotations

import numpy as np

AUG_FLIP_H, AUG_FLIP_W, AUG_TRANSPOSE = 1, 2, 4


def preprocess_tail(raw: np.ndarray, norm_fac: float | None, log_scale: bool) -> np.ndarray:
    """dataset.py:215-222: astype(float32); log(max(x, 1e-10)); /= norm_fac."""
    x = raw.astype(np.float32)
    if log_scale:
        x = np.log(np.maximum(x, 1e-10))
    if norm_fac is not None:
        x /= norm_fac
    return x


def transform_rasters(x: np.ndarray, bits: int) -> np.ndarray:
    """dataset.py:232-256 on one raster [D, C, H, W]: flip axis 2, flip axis 3, swapaxes(2, 3), contiguous."""
    if bits & AUG_FLIP_H:
        x = np.flip(x, axis=2)
    if bits & AUG_FLIP_W:
        x = np.flip(x, axis=3)
    if bits & AUG_TRANSPOSE:
        x = np.swapaxes(x, 2, 3)
    return np.ascontiguousarray(x)


def raster_pipeline(raw: np.ndarray, aug: np.ndarray | None, norm_fac: float | None, log_scale: bool) -> np.ndarray:
    """Batch form: raw [B, D, C, H, W] on-disk dtype, aug [B] bits -> fp32 [B, D, C, H', W']."""
    out = []
    for b in range(raw.shape[0]):
        out.append(transform_rasters(preprocess_tail(raw[b], norm_fac, log_scale), int(aug[b]) if aug is not None else 0))
    return np.stack(out)


# ------------------------------------------------------------------ Philox4x32-10 (Salmon et al., SC'11)
_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = 0x9E3779B9, 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(ctr: np.ndarray, key: np.ndarray) -> np.ndarray:
    """ctr [..., 4] uint32, key [..., 2] uint32 (broadcastable) -> [..., 4] uint32."""
    c = [np.asarray(ctr[..., i], dtype=np.uint64) for i in range(4)]
    k0 = np.asarray(key[..., 0], dtype=np.uint64)
    k1 = np.asarray(key[..., 1], dtype=np.uint64)
    for _ in range(10):
        p0, p1 = _M0 * c[0], _M1 * c[2]
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & _MASK, p1 >> np.uint64(32), p1 & _MASK
        c = [hi1 ^ c[1] ^ k0, lo1, hi0 ^ c[3] ^ k1, lo0]
        k0 = (k0 + np.uint64(_W0)) & _MASK
        k1 = (k1 + np.uint64(_W1)) & _MASK
    return np.stack([x.astype(np.uint32) for x in np.broadcast_arrays(*c)], axis=-1)


def _uniform(seed: int, offset: int, sample: int, pas: int, mod: int, kind: int, index: np.ndarray) -> np.ndarray:
    """The device kernel's draw: counter (index, kind | mod << 8, sample, lo32(offset + pass)), 24-bit mantissa."""
    index = np.asarray(index, dtype=np.uint32)
    n = (offset + pas) & (2**64 - 1)
    ctr = np.zeros(index.shape + (4,), dtype=np.uint32)
    ctr[..., 0] = index
    ctr[..., 1] = kind | (mod << 8)
    ctr[..., 2] = sample
    ctr[..., 3] = n & 0xFFFFFFFF
    key = np.array([seed & 0xFFFFFFFF, ((seed >> 32) ^ (n >> 32)) & 0xFFFFFFFF], dtype=np.uint32)
    x = philox4x32_10(ctr, key)[..., 0]
    return (x >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)


def struct_mask(mods: list[tuple], seed: int, offset: int, B: int, max_pass: int = 1 << 20):
    """mods: [(G, D, L, p_mod, p_bands, p_dates, p_loc)] of one group -> (mask [B, L_group] uint8, passes [B])."""
    Lg = sum(G * D * L for G, D, L, *_ in mods)
    out = np.zeros((B, Lg), dtype=np.uint8)
    passes = np.zeros(B, dtype=np.int32)
    for b in range(B):
        pas = 0
        while True:
            parts = []
            for m, (G, D, L, pm, pb, pd, pl) in enumerate(mods):
                acc = np.zeros((G, D, L), dtype=bool)        # mae.py:195-198
                if pm and pm > 0:                            # :199-200
                    acc |= _uniform(seed, offset, b, pas, m, 0, np.zeros(1))[0] < np.float32(pm)
                if pb and pb > 0:                            # :201-204
                    acc |= (_uniform(seed, offset, b, pas, m, 1, np.arange(G)) < np.float32(pb))[:, None, None]
                if pd and pd > 0:                            # :205-208
                    acc |= (_uniform(seed, offset, b, pas, m, 2, np.arange(D)) < np.float32(pd))[None, :, None]
                if pl and pl > 0:                            # :209-210
                    acc |= (_uniform(seed, offset, b, pas, m, 3, np.arange(L)) < np.float32(pl))[None, None, :]
                parts.append(acc.reshape(-1))                # :211 reshape (B, D, L, 1), grouped in order
            row = np.concatenate(parts)
            pas += 1
            if not row.all() or pas >= max_pass:             # :190-193, 216-221: redraw while entirely masked
                break
        out[b], passes[b] = row, pas
    return out, passes
