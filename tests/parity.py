"""Comparator used by the GPU parity tests and by __graft_entry__.smoke().

Tolerances are BASELINE.json's north_star:
  * hit mask and object id exact, except on stated grazing-angle pixels (counted and bounded);
  * position and normal within 1e-4 relative (on pixels whose object ids agree);
  * linear HDR colour within 1e-3 absolute per channel on >= 99.9 % of pixels.
"""
from __future__ import annotations

import numpy as np

COLOR_ABS_TOL = 1e-3
GEOM_REL_TOL = 1e-4
NORMAL_FP16_TOL = 1.0e-3      # gNormal is RGBA16F in the reference (4.9e-4 relative per component)


def compare(kernel: dict, oracle: dict, normal_fp16: bool = False) -> dict:
    kid, oid = kernel["object_id"], oracle["object_id"]
    n = kid.size
    hit_mismatch = int(((kid >= 0) != (oid >= 0)).sum())
    id_mismatch = int((kid != oid).sum())
    same = kid == oid
    kc, oc = kernel["color"][..., :3], oracle["color"][..., :3]
    both_nan = np.isnan(kc) & np.isnan(oc)
    both_inf = np.isinf(kc) & np.isinf(oc) & (np.sign(kc) == np.sign(oc))
    with np.errstate(invalid="ignore"):
        dc = np.abs(kc - oc)
    dc = np.where(both_nan | both_inf, 0.0, dc)
    dc = np.where(np.isnan(dc), np.inf, dc)
    bad_color = (dc > COLOR_ABS_TOL).any(axis=-1)
    kp, op = kernel["position"][..., :3], oracle["position"][..., :3]
    scale_p = np.maximum(np.linalg.norm(op, axis=-1), 1.0)
    with np.errstate(invalid="ignore"):
        dp = np.linalg.norm(kp - op, axis=-1) / scale_p
    kn, on = kernel["normal"][..., :3], oracle["normal"][..., :3]
    with np.errstate(invalid="ignore"):
        dn = np.abs(kn - on).max(axis=-1)
    dp = np.where(np.isnan(dp), np.inf, dp)
    dn = np.where(np.isnan(dn), np.inf, dn)
    ntol = NORMAL_FP16_TOL if normal_fp16 else GEOM_REL_TOL
    bad_p = same & (dp > GEOM_REL_TOL)
    bad_n = same & (dn > ntol)
    finite = np.isfinite(dc).all(axis=-1)
    return {
        "pixels": n,
        "hit_mismatch": hit_mismatch,
        "id_mismatch": id_mismatch,
        "color_bad": int(bad_color.sum()),
        "color_bad_frac": float(bad_color.mean()),
        "color_bad_same_id": int((bad_color & same).sum()),
        "color_max_abs_same_id": float(dc[same & finite].max()) if (same & finite).any() else 0.0,
        "position_bad": int(bad_p.sum()),
        "normal_bad": int(bad_n.sum()),
        "position_max_rel_same_id": float(dp[same].max()) if same.any() else 0.0,
        "normal_max_abs_same_id": float(dn[same].max()) if same.any() else 0.0,
        "alpha_ok": bool((kernel["color"][..., 3] == 1.0).all() and (kernel["position"][..., 3] == 1.0).all()),
    }


def assert_parity(stats: dict, max_id_frac: float = 2e-3, max_color_frac: float = 1e-3, max_geom_frac: float = 1e-3):
    n = stats["pixels"]
    assert stats["alpha_ok"], stats
    assert stats["id_mismatch"] <= max(2, max_id_frac * n), ("object-id mismatches beyond the grazing-pixel bound", stats)
    assert stats["hit_mismatch"] <= max(2, max_id_frac * n), ("hit-mask mismatches beyond the grazing-pixel bound", stats)
    assert stats["color_bad_frac"] <= max_color_frac, ("colour outside 1e-3 abs on more than 0.1% of pixels", stats)
    assert stats["position_bad"] <= max(2, max_geom_frac * n), ("position outside 1e-4 rel", stats)
    assert stats["normal_bad"] <= max(2, max_geom_frac * n), ("normal outside tolerance", stats)
