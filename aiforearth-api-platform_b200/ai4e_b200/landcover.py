"""Land-cover pipeline (BASELINE.json configs[3]): one large multi-band tile -> non-overlapping
chips -> per-chip classifier.  Chips are never copied on the host: each one is submitted as a
strided view of the tile (row stride = tile width x bands) and the engine's staging copy does the
gather.  Chips are independent, so with a `Pool` they spread over the GPUs with no halo exchange."""
from __future__ import annotations

import numpy as np


def chip_views(tile_u8_hwc: np.ndarray, chip: int = 256):
    """yield ((row, col), view) for every full chip of the tile, row-major over the chip grid."""
    h, w, _ = tile_u8_hwc.shape
    for r in range(h // chip):
        for c in range(w // chip):
            yield (r, c), tile_u8_hwc[r * chip:(r + 1) * chip, c * chip:(c + 1) * chip]


def classify_tile(engine, tile_u8_hwc: np.ndarray, chip: int = 256):
    """-> (class_id [gh, gw] int32, top5_idx [gh, gw, 5] int32, top5_prob [gh, gw, 5] float32).

    `engine` is an `Engine` or a `Pool`; its model must take `chip x chip` inputs with the tile's
    band count."""
    a = np.asarray(tile_u8_hwc)
    if a.dtype != np.uint8 or a.ndim != 3:
        raise ValueError("expected a uint8 [H, W, bands] tile")
    gh, gw = a.shape[0] // chip, a.shape[1] // chip
    if hasattr(engine, "submit_tile"):  # one C call for the whole chip grid, one for the results
        tickets = engine.submit_tile(a, chip)
        engine.flush()
        idx, prob = engine.wait_many(tickets)
        idx, prob = idx.reshape(gh, gw, 5), prob.reshape(gh, gw, 5)
        return idx[:, :, 0].copy(), idx, prob
    tickets = [(rc, engine.submit(v)) for rc, v in chip_views(a, chip)]
    engine.flush()
    idx = np.empty((gh, gw, 5), np.int32)
    prob = np.empty((gh, gw, 5), np.float32)
    for (r, c), t in tickets:
        idx[r, c], prob[r, c] = engine.wait(t)
    return idx[:, :, 0].copy(), idx, prob


def summarize_tile(engine, tile_u8_hwc: np.ndarray, chip: int = 256) -> dict:
    """tile -> chips -> classify -> aggregate in one submission, the aggregate computed on the GPU(s) that
    classified the chips (K7): the tile's label map, its class histogram and the per-class probability mass.
    The reference would chain the aggregation as one more task (`AddPipelineTask`,
    Common/task_management/distributed_api_task.py:67-100): one HTTP hop and one task-store write per stage.
    Keys: labels [gh, gw], label_prob [gh, gw], hist [classes], mass_q32 / mass [classes]."""
    ticket, grid = engine.submit_tile_summary(np.asarray(tile_u8_hwc), chip)
    return engine.wait_tile_summary(ticket, grid)
