"""Helpers shared by the tests: oracle model -> engine program, golden loading."""
import functools
import os

import numpy as np

from oracle import preprocess_oracle as pp
from oracle import resnet_oracle as ro

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@functools.lru_cache(maxsize=4)
def oracle_model(arch: str, in_ch: int = 3, hw: int = 224):
    return ro.build_model(arch, in_ch=in_ch, calib_hw=hw)


@functools.lru_cache(maxsize=4)
def program_for(arch: str, in_ch: int = 3, hw: int = 224, keep_stages: bool = False):
    from ai4e_b200.models import build_resnet_program
    m = oracle_model(arch, in_ch, hw)
    return build_resnet_program(arch, m.state_dict(), hw, hw, keep_stages=keep_stages)


def load_golden(name: str):
    g = np.load(os.path.join(GOLDEN, f"classify_{name}.npz"))
    hw, in_ch = int(g["hw"]), int(g["in_ch"])
    imgs = ro.synthetic_images([int(i) for i in g["indices"]], hw, hw, in_ch, seed=int(g["image_seed"]))
    return g, imgs


def mean_std(in_ch: int):
    return (ro.MEAN4, ro.STD4) if in_ch == 4 else (ro.IMAGENET_MEAN, ro.IMAGENET_STD)


# north_star tolerance for floating point: |got - ref| <= 1e-2 + 1e-3 |ref|
ATOL, RTOL = 1e-2, 1e-3


def within_north_star(got: np.ndarray, ref: np.ndarray) -> np.ndarray:
    return np.abs(got - ref) <= ATOL + RTOL * np.abs(ref)
