"""Pins the tiling order, the LRMS up-sampling and the quality-metric restatement against the UNMODIFIED reference
(data/LRHR_dataset.py patch_16 / unpatch_16 / Upsample, core/metrics.py SAM / ERGAS / UIQC) and stores small vectors.

    python tests/golden/make_golden_scene.py      (build container only: needs /root/reference)
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(HERE, "_shims"), "/root/reference", ROOT]
for name in ("skimage", "skimage.metrics", "sewar", "sewar.full_ref", "h5py"):       # imported at module top, unused here
    sys.modules.setdefault(name, types.ModuleType(name))
sys.modules["skimage.metrics"].peak_signal_noise_ratio = sys.modules["skimage.metrics"].structural_similarity = None
import scipy.ndimage  # noqa: E402
sys.modules.setdefault("scipy.ndimage.filters", scipy.ndimage)

from oracle import metrics_ref  # noqa: E402


def main():
    import core.metrics as M
    import data.LRHR_dataset as L
    g = np.random.default_rng(5)
    out = {}
    tiles = g.random((16, 3, 8, 8)).astype(np.float32)
    cube = L.patch_16(tiles)
    assert np.array_equal(L.unpatch_16(cube.astype(np.float32)), tiles)
    out["tiles16"], out["cube16"] = tiles, cube.astype(np.float32)
    lr = g.random((2, 4, 16, 16)).astype(np.float32)
    out["lrms"], out["lms"] = lr, L.Upsample(lr, 64).astype(np.float32)
    a = g.random((32, 32, 4)) * 0.8 + 0.1
    b = np.clip(a + g.normal(0, 0.02, a.shape), 0.01, 1)
    ref = {"sam": M.SAM_numpy(a, b), "ergas": M.ERGAS_numpy(a, b), "uiqc": M.UIQC_numpy(a, b)}
    mine = {"sam": metrics_ref.sam(a, b), "ergas": metrics_ref.ergas(a, b), "uiqc": metrics_ref.uiqc(a, b)}
    print(ref, mine)
    assert all(abs(ref[k] - mine[k]) <= 1e-12 * abs(ref[k]) for k in ref), (ref, mine)
    out["metric_a"], out["metric_b"] = a, b
    out["metric_values"] = np.array([ref["sam"], ref["ergas"], ref["uiqc"]])
    np.savez_compressed(os.path.join(HERE, "scene_vectors.npz"), **out)


if __name__ == "__main__":
    main()
