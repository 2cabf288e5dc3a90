"""Similarity -> rigid split used for object poses.

Mirrors gsworld/utils/pcd_utils.py:224-252 (extract_rigid_transform): SVD of the 3x3 block,
scale = mean singular value, R = U @ Vh (no reflection fix), translation kept.
"""
from __future__ import annotations

import torch


def extract_rigid_transform(M: torch.Tensor):
    if M.ndim == 2 and M.shape == (4, 4):
        M = M.unsqueeze(0)
        single = True
    elif M.ndim == 3 and M.shape[1:] == (4, 4):
        single = False
    else:
        raise ValueError("Input must be (4,4) or (N,4,4)")
    A, t = M[:, :3, :3], M[:, :3, 3]
    U, S, Vh = torch.linalg.svd(A)
    scales = S.mean(dim=1)
    R = torch.matmul(U, Vh)
    out = torch.eye(4, device=M.device, dtype=M.dtype).repeat(M.shape[0], 1, 1)
    out[:, :3, :3] = R
    out[:, :3, 3] = t
    if single:
        return out[0], scales[0], R[0], t[0]
    return out, scales, R, t


def pose7_to_matrix(pose: torch.Tensor) -> torch.Tensor:
    """(B,7) p + wxyz quaternion -> (B,4,4), as ManiSkill's Pose.create(...).to_transformation_matrix()
    (call site gs_world_wrapper.py:139-140; external, restated)."""
    p, q = pose[:, :3], pose[:, 3:7]
    q = q / q.norm(dim=-1, keepdim=True)
    w, x, y, z = q.unbind(-1)
    R = torch.stack(
        [
            1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
            2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
            2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y),
        ],
        dim=-1,
    ).reshape(-1, 3, 3)
    M = torch.eye(4, dtype=pose.dtype, device=pose.device).repeat(pose.shape[0], 1, 1)
    M[:, :3, :3] = R
    M[:, :3, 3] = p
    return M
