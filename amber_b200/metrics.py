"""Image-parity metrics: luminance, 4x4 box downsample, relative RMSE.

The reference publishes no tolerance; its implicit criterion is that its algorithms converge to the
same image (``/root/reference/README.md``).  The figures used here:

  rRMSE(I, R) = sqrt(mean((I - R)^2)) / mean(R)   on luminance after a 4x4 box downsample,

evaluated separately above and below the Cornell box's water line (the sheet at y = -0.5,
``src/amber/etude/cornel_box.cc:156-176``): below it every light path is a caustic that only BSDF
sampling finds, so both estimators are equally noisy there.
"""
from __future__ import annotations

import numpy as np


def luminance(image: np.ndarray) -> np.ndarray:
    """Rec. 709 luminance of an (H, W, 3) image, float64."""
    image = np.asarray(image, dtype=np.float64)
    return 0.2126 * image[..., 0] + 0.7152 * image[..., 1] + 0.0722 * image[..., 2]


def box_downsample(plane: np.ndarray, k: int = 4) -> np.ndarray:
    h, w = plane.shape
    return plane[: h - h % k, : w - w % k].reshape(h // k, k, w // k, k).mean(axis=(1, 3))


def rrmse(image: np.ndarray, reference: np.ndarray, rows: slice | None = None, k: int = 4) -> float:
    """Relative RMSE of the k x k block means of luminance; ``rows`` selects block rows."""
    a, b = box_downsample(luminance(image), k), box_downsample(luminance(reference), k)
    if rows is not None:
        a, b = a[rows], b[rows]
    return float(np.sqrt(np.mean((a - b) ** 2)) / max(np.mean(b), 1e-300))


def water_line_rows(height: int, first_row: int = 0, n_rows: int | None = None, k: int = 4) -> tuple[slice, slice]:
    """Block-row slices (after a k x k downsample of rows [first_row, first_row + n_rows) of a
    ``height``-row frame) that lie entirely above / entirely below the water of the built-in Cornell
    box.  The thin lens at z = 4 (sensor distance 0.0506, 36 mm film) sees the sheet y = -0.5 between
    image rows 0.64 H (at the back wall) and 0.74 H (at the front edge); rows above 0.62 H see no water,
    rows below 0.76 H only see it (or the floor through it)."""
    n_rows = height - first_row if n_rows is None else n_rows
    above_end = int(np.floor((0.62 * height - first_row) / k))
    below_begin = int(np.ceil((0.76 * height - first_row) / k))
    n_blocks = n_rows // k
    above = slice(0, max(0, min(n_blocks, above_end)))
    below = slice(max(0, min(n_blocks, below_begin)), n_blocks)
    return above, below
