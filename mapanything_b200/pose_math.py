"""Host-side pose algebra of the optional camera-pose inputs (a handful of floats per view).

The reference does this on [B,4]/[B,3] tensors before its pose encoders run
(mapanything/models/mapanything/model.py:792-896, mapanything/utils/geometry.py:601-852,1629-1666).
It is control-plane sized work (views x 7 numbers), kept above the C ABI: `mab_inputs` takes the
quaternions / translations already relative to view 0 plus the scene's translation norm.
All functions work on CPU float32 tensors with quaternions in xyzw order.
"""
import torch


def quat_to_rotmat(q):
    """geometry.py:601-652: re-normalises, returns (N,3,3)."""
    q = q / q.norm(dim=1, keepdim=True)
    x, y, z, w = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    rows = [1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
            2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
            2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]
    return torch.stack(rows, dim=1).view(-1, 3, 3)


def rotmat_to_quat(m):
    """geometry.py:655-713: the four-candidate construction, pick the best-conditioned one,
    return xyzw with w >= 0."""
    m = m.reshape(-1, 3, 3).float()
    m00, m01, m02 = m[:, 0, 0], m[:, 0, 1], m[:, 0, 2]
    m10, m11, m12 = m[:, 1, 0], m[:, 1, 1], m[:, 1, 2]
    m20, m21, m22 = m[:, 2, 0], m[:, 2, 1], m[:, 2, 2]
    sq = torch.stack([1.0 + m00 + m11 + m22, 1.0 + m00 - m11 - m22,
                      1.0 - m00 + m11 - m22, 1.0 - m00 - m11 + m22], dim=-1)
    q_abs = torch.where(sq > 0, torch.sqrt(sq.clamp(min=0)), torch.zeros_like(sq))
    # candidate r (wxyz) is exact when component r is the largest one
    cands = torch.stack([
        torch.stack([q_abs[:, 0] ** 2, m21 - m12, m02 - m20, m10 - m01], dim=-1),
        torch.stack([m21 - m12, q_abs[:, 1] ** 2, m10 + m01, m02 + m20], dim=-1),
        torch.stack([m02 - m20, m10 + m01, q_abs[:, 2] ** 2, m12 + m21], dim=-1),
        torch.stack([m10 - m01, m20 + m02, m21 + m12, q_abs[:, 3] ** 2], dim=-1)], dim=-2)
    cands = cands / (2.0 * q_abs[:, :, None].clamp(min=0.1))
    best = q_abs.argmax(dim=-1)
    wxyz = cands[torch.arange(m.shape[0]), best]
    xyzw = wxyz[:, [1, 2, 3, 0]]
    return torch.where(xyzw[:, 3:4] < 0, -xyzw, xyzw)


def quat_mul(a, b):
    """geometry.py:774-808 (Hamilton product, xyzw)."""
    x1, y1, z1, w1 = a[:, 0], a[:, 1], a[:, 2], a[:, 3]
    x2, y2, z2, w2 = b[:, 0], b[:, 1], b[:, 2], b[:, 3]
    return torch.stack([w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                        w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                        w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2,
                        w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2], dim=1)


def relative_to_first(q_ref, t_ref, q, t):
    """geometry.py:744-772,814-852: poses (q,t) expressed in the frame of (q_ref,t_ref)."""
    n = q.shape[0]
    q_ref, t_ref = q_ref.expand(n, 4), t_ref.expand(n, 3)
    conj = torch.cat([-q_ref[:, :3], q_ref[:, 3:]], dim=1)
    q_inv = conj / (q_ref * q_ref).sum(dim=1, keepdim=True)
    r_inv = quat_to_rotmat(q_inv)
    t_inv = -torch.einsum("bij,bj->bi", r_inv, t_ref)
    return quat_mul(q_inv, q), torch.einsum("bij,bj->bi", r_inv, t) + t_inv


def translation_norm(trans):
    """geometry.py:1629-1666: mean length of the non-zero translations of the scene, >= 1e-8."""
    dis = trans.norm(dim=-1)
    return float((dis.sum() / ((dis > 0).sum() + 1e-8)).clip(min=1e-8))


def split_camera_poses(cp, view_idx=0):
    """inference.py:276-296: (quats, trans) tuple or (B,4,4) cam2world matrices -> (quats, trans)."""
    if isinstance(cp, tuple) and len(cp) == 2:
        return cp[0], cp[1]
    if torch.is_tensor(cp) and cp.shape[-2:] == (4, 4):
        cp = cp.detach().float().cpu()
        return rotmat_to_quat(cp[:, :3, :3]), cp[:, :3, 3]
    raise ValueError(f"View {view_idx}: camera_poses must be either a tuple of (quats, trans) "
                     f"or a tensor of (B, 4, 4) transformation matrices.")
