"""Drop-in ``PinholeCamera`` / ``SDFMarcher`` / ``SDFNormals`` (reference ``rendering/ray_marching.py``).

Same constructor arguments, ``forward`` signatures and return shapes; bodies call the CUDA
kernels.  When autograd is recording, the march and every SDF evaluation go through custom ops
whose backward is the replay kernel in ``csrc/rm_backward.cu``; the thin elementwise glue around
them (camera rotation, stencil algebra) stays in torch so that ``torch.autograd`` composes.
"""
from __future__ import annotations

import torch
import torch.nn as nn
from torch import Tensor

from .. import _lib, ops
from .. import quaternion as Q
from ..tape import tape_of


def _grad_mode(*tensors) -> bool:
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors)


class PinholeCamera(nn.Module):
    """Sensor-plane pinhole camera (reference ``ray_marching.py:9-64``).

    Rays start on the sensor plane (z = 0 in the camera frame) and point away from the focal
    point ``(0, 0, -focal_length)``; row 0 is +y, the camera looks down +z."""

    def __init__(self, num_cameras: int, px_width: int, px_height: int, focal_length: float,
                 sensor_width: float, sensor_height: float):
        super().__init__()
        self.num_cameras = num_cameras
        self.focal_length = focal_length
        self.sensor_width = sensor_width
        self.sensor_height = sensor_height
        self.size = (num_cameras, 1, px_height, px_width)
        self.px_width, self.px_height = px_width, px_height
        # the reference's small buffers (ray_marching.py:26-39,51-54), same names / shapes (stored
        # contiguous: the reference's expanded views cannot be written by load_state_dict for N > 1)
        self.register_buffer("focus", torch.tensor([[[[0., 0., -focal_length]]]]).expand(num_cameras, 1, 1, 3).clone())
        self.register_buffer("theta", torch.tensor([[[sensor_width / 2, 0., 0.],
                                                     [0., -sensor_height / 2, 0.]]]).expand(num_cameras, 2, 3).clone())
        self.register_buffer("pixel_frames",
                             torch.eye(3)[None, None, None, [0, 1], :].expand(num_cameras, 1, 1, 2, 3).clone())
        self.quaternion_to_so3 = Q.QuaternionToSO3()
        self._base = None
        # `ray_positions` / `ray_directions` (ray_marching.py:40-50) are 24 B/ray of state the kernels
        # never read (rays come from the pixel index): they are not held as buffers, but state_dict()
        # carries them (materialised on demand) and load_state_dict() accepts them, so checkpoints
        # move between the reference and this module with strict=True in both directions.
        self._register_state_dict_hook(PinholeCamera._emit_ray_buffers)
        self._register_load_state_dict_pre_hook(self._accept_ray_buffers)

    def struct(self):
        return ops.camera_struct(self.num_cameras, self.px_width, self.px_height,
                                 self.focal_length, self.sensor_width, self.sensor_height)

    def closed_form_rays(self, device=None):
        """The reference's ``ray_positions`` / ``ray_directions`` buffers from their closed form
        (SURVEY.md 8(a) C1), in plain torch on any device: pixel (row i, col j) ->
        ``(sw/2 ((2j+1)/W - 1), -sh/2 ((2i+1)/H - 1), 0)``, ``dir = normalize(pos - focus)``."""
        N, H, W = self.num_cameras, self.px_height, self.px_width
        device = device if device is not None else self.focus.device
        gx = (2 * torch.arange(W, device=device, dtype=torch.float32) + 1) / W - 1
        gy = (2 * torch.arange(H, device=device, dtype=torch.float32) + 1) / H - 1
        x = (self.sensor_width / 2 * gx)[None, :].expand(H, W)
        y = (-self.sensor_height / 2 * gy)[:, None].expand(H, W)
        pos = torch.stack([x, y, torch.zeros_like(x)], dim=-1)[None].expand(N, H, W, 3).contiguous()
        rel = pos - self.focus.to(device)
        return pos, rel / rel.norm(dim=-1, keepdim=True)

    @staticmethod
    def _emit_ray_buffers(module, state_dict, prefix, local_metadata):
        pos, dirs = module.closed_form_rays()
        # keep the reference's key order: focus, theta, ray_positions, ray_directions, pixel_frames, ...
        items = list(state_dict.items())
        state_dict.clear()
        for k, v in items:
            if k == prefix + "pixel_frames":
                state_dict[prefix + "ray_positions"] = pos
                state_dict[prefix + "ray_directions"] = dirs
            state_dict[k] = v
        return state_dict

    def _accept_ray_buffers(self, state_dict, prefix, local_metadata, strict, missing_keys,
                            unexpected_keys, error_msgs):
        shape = (self.num_cameras, self.px_height, self.px_width, 3)
        for name in ("ray_positions", "ray_directions"):
            t = state_dict.pop(prefix + name, None)
            if t is None:
                if strict:
                    missing_keys.append(prefix + name)
            elif tuple(t.shape) != shape:
                error_msgs.append(f"size mismatch for {prefix + name}: copying a param with shape "
                                  f"{tuple(t.shape)} from checkpoint, the shape in current model is {shape}.")

    def base_rays(self, device):
        """Camera-frame rays (the reference's ``ray_positions`` / ``ray_directions`` buffers),
        generated on the device by the ray kernel with an identity pose and cached."""
        if self._base is None or self._base[0].device != device:
            ident = torch.tensor([[1.0, 0.0, 0.0, 0.0]], device=device).expand(self.num_cameras, 4)
            zero = torch.zeros(self.num_cameras, 3, device=device)
            with torch.no_grad():
                o, v = ops.camera_rays(self.struct(), ops.pose_records(ident.contiguous(), zero))
            shape = (self.num_cameras, self.px_height, self.px_width, 3)
            self._base = (o.reshape(shape), v.reshape(shape))
        return self._base

    @property
    def ray_positions(self):
        return self.base_rays(torch.device("cuda", torch.cuda.current_device()))[0]

    @property
    def ray_directions(self):
        return self.base_rays(torch.device("cuda", torch.cuda.current_device()))[1]

    def forward(self, orientation: Tensor, translation: Tensor):
        dev = _lib.require_cuda(orientation, translation)
        shape = (self.num_cameras, self.px_height, self.px_width, 3)
        frames = self.quaternion_to_so3(orientation[:, None, None, :])
        if _grad_mode(orientation, translation):
            o, v = self.base_rays(dev)
            R = Q.rotation_matrix(orientation)[:, None, None]            # (N,1,1,3,3)
            pos = (R @ o[..., None])[..., 0] + translation[:, None, None, :]
            dirs = (R @ v[..., None])[..., 0]
        else:
            with torch.no_grad():
                pos, dirs = ops.camera_rays(self.struct(), ops.pose_records(orientation, translation))
            pos, dirs = pos.reshape(shape), dirs.reshape(shape)
        return (pos, frames, pos, dirs)


class SDFMarcher(nn.Module):
    """Fixed-count sphere tracer (reference ``ray_marching.py:67-84``)."""

    def __init__(self, sdf_scene: nn.Module):
        super().__init__()
        self.sdf_scene = sdf_scene

    def forward(self, ray_positions: Tensor, ray_directions: Tensor,
                marching_steps: int = 32) -> Tensor:
        return ops.march(tape_of(self.sdf_scene), ray_positions, ray_directions, marching_steps)


class SDFNormals(nn.Module):
    """Tetrahedral-stencil normals and Laplacian (reference ``ray_marching.py:87-125``)."""

    def __init__(self, sdf_scene: nn.Module, normals_eps: float = 1e-3):
        super().__init__()
        self.sdf_scene = sdf_scene
        self.normals_eps = normals_eps
        s = 0.5 ** 0.5
        off = torch.tensor([[1.0, 0.0, -s], [-1.0, 0.0, -s], [0.0, 1.0, s], [0.0, -1.0, s]])
        off = off / off.norm(dim=-1, keepdim=True) * normals_eps
        rel = off[1:] - off[:1]
        self.register_buffer("offsets", off)
        self.register_buffer("relative_offsets", rel)
        self.register_buffer("offsets_inverse", torch.linalg.inv(rel))

    def forward(self, surface_coords: Tensor):
        tape = tape_of(self.sdf_scene)
        if not (_grad_mode(surface_coords) or _grad_mode(*tape.params)):
            return ops.normals(tape, surface_coords, self.normals_eps)
        # differentiable route: five tape evaluations through the sdf_eval op, stencil in torch
        samples = ops.sdf_eval(tape, surface_coords[..., None, :] + self.offsets)[..., 0]  # (...,4)
        delta = samples[..., 1:] - samples[..., :1]
        n = delta @ self.offsets_inverse.T
        n = n / n.norm(dim=-1, keepdim=True)
        centre = ops.sdf_eval(tape, surface_coords)
        lap = (centre - samples.mean(dim=-1, keepdim=True)) * (6 / self.normals_eps ** 2)
        return (n, lap)
