"""Host-side pieces of ray_marching_b200.frames (no GPU): the PNG / NPY writers and argument checks."""
import numpy as np
import pytest
import torch

from ray_marching_b200 import frames
from ray_marching_b200._lib import RmError


def test_png_round_trip(tmp_path):
    g = torch.Generator().manual_seed(3)
    img = torch.rand(37, 53, 4, generator=g)
    img[0, 0] = torch.tensor([float("nan"), -1.0, 2.0, 1.0])      # NaN -> 0, clamp to [0, 1]
    path = tmp_path / "frame.png"
    frames.write_png(path, img)
    back = frames.read_png(path)
    want = frames.to_rgba8(img)
    assert back.shape == (37, 53, 4) and back.dtype == np.uint8
    assert (back == want).all()
    assert tuple(back[0, 0]) == (0, 0, 255, 255)
    frames.write_png(tmp_path / "rgb.png", img[..., :3])
    assert (frames.read_png(tmp_path / "rgb.png") == want[..., :3]).all()
    with pytest.raises(ValueError):
        frames.write_png(tmp_path / "bad.png", img[..., :2])


def test_npy_dump(tmp_path):
    x = torch.arange(24, dtype=torch.float32).reshape(2, 3, 4)
    frames.write_npy(tmp_path / "x.npy", x)
    assert (np.load(tmp_path / "x.npy") == x.numpy()).all()


def test_camera_path_has_no_cpu_fallback():
    with pytest.raises(RmError):
        frames.CameraPath([(0.0, 0.0, 1.0)], [(1.0, 0.0, 0.0, 0.0)], device="cpu")
