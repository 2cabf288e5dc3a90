"""Camera matrices in the convention the rasterizer consumes.

Mirrors submodules/gaussian-splatting/scene/cameras.py:19-89 (Camera) and
utils/graphics_utils.py:38-71 (getWorld2View2, getProjectionMatrix): all three tensors are the
TRANSPOSED (row-vector convention) matrices; znear 0.01, zfar 100.  The PIL image / alpha-mask
plumbing of the reference class is training-side and is not carried.
"""
from __future__ import annotations

import math

import numpy as np
import torch


def getWorld2View2(R, t, translate=np.array([0.0, 0.0, 0.0]), scale=1.0):
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.transpose()
    Rt[:3, 3] = t
    Rt[3, 3] = 1.0
    C2W = np.linalg.inv(Rt)
    C2W[:3, 3] = (C2W[:3, 3] + translate) * scale
    return np.float32(np.linalg.inv(C2W))


def getProjectionMatrix(znear, zfar, fovX, fovY):
    tanHalfFovY = math.tan(fovY / 2)
    tanHalfFovX = math.tan(fovX / 2)
    top, right = tanHalfFovY * znear, tanHalfFovX * znear
    bottom, left = -top, -right
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


class Camera:
    """Duck-type of the reference Camera: the attributes render() reads
    (gaussian_renderer/__init__.py:33-46): FoVx, FoVy, image_height, image_width,
    world_view_transform, full_proj_transform, camera_center, image_name."""

    def __init__(self, resolution, colmap_id, R, T, FoVx, FoVy, depth_params=None, image=None,
                 invdepthmap=None, image_name="", uid=0, trans=np.array([0.0, 0.0, 0.0]), scale=1.0,
                 data_device="cuda", **_unused):
        self.uid, self.colmap_id = uid, colmap_id
        self.R, self.T = R, T
        self.FoVx, self.FoVy = FoVx, FoVy
        self.image_name = image_name
        self.data_device = torch.device(data_device)
        self.image_width, self.image_height = int(resolution[0]), int(resolution[1])
        self.zfar, self.znear = 100.0, 0.01
        self.trans, self.scale = trans, scale
        dev = self.data_device
        self.world_view_transform = torch.tensor(getWorld2View2(R, T, trans, scale)).transpose(0, 1).to(dev)
        self.projection_matrix = getProjectionMatrix(self.znear, self.zfar, FoVx, FoVy).transpose(0, 1).to(dev)
        self.full_proj_transform = (
            self.world_view_transform.unsqueeze(0).bmm(self.projection_matrix.unsqueeze(0))
        ).squeeze(0)
        self.camera_center = self.world_view_transform.inverse()[3, :3]
