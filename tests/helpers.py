"""Shared comparison helpers for the parity tests (golden fixtures <-> oracle <-> CUDA path)."""
import math
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

MASKED_KEYS = ("pts3d", "pts3d_cam", "depth_along_ray", "depth_z")


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN_DIR, name + ".npz")))


def unpack_mask(g, key):
    shape = tuple(int(x) for x in g[key + ".shape"])
    return np.unpackbits(g[key + ".bits"])[: int(np.prod(shape))].astype(bool).reshape(shape)


def rel_l2(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))


def quat_angle_deg(q, q_ref):
    """Angular error between unit quaternions, 4*asin(|q - q_ref|/2) after sign alignment
    (mapanything/utils/metrics.py:18-34)."""
    q = np.asarray(q, np.float64).reshape(-1, 4)
    r = np.asarray(q_ref, np.float64).reshape(-1, 4)
    q = q / np.linalg.norm(q, axis=1, keepdims=True)
    r = r / np.linalg.norm(r, axis=1, keepdims=True)
    sign = np.where((q * r).sum(1, keepdims=True) < 0, -1.0, 1.0)
    d = np.linalg.norm(q * sign - r, axis=1)
    return float(np.degrees(4 * np.arcsin(np.clip(d / 2, 0, 1))).max())


def compare_with_golden(preds, g, stride, tol_dense, tol_pose_deg, tol_small=None, min_mask_agree=0.0):
    """preds: list of per-view dicts (torch tensors, any device).  Returns a report dict and asserts.

    Dense float keys are compared by relative L2 on the golden's sub-sampled grid; keys that
    `infer` multiplies by the final mask are compared where both masks are True (the mask itself is
    compared by agreement rate: it thresholds floats produced by different arithmetic)."""
    report = {}
    nviews = len(preds)
    for i in range(nviews):
        p = {k: (v.detach().float().cpu().numpy() if torch.is_tensor(v) and v.dtype != torch.bool
                 else (v.detach().cpu().numpy() if torch.is_tensor(v) else v)) for k, v in preds[i].items()}
        gm = unpack_mask(g, f"v{i}.mask")
        pm = p["mask"].astype(bool)
        assert pm.shape == gm.shape, (pm.shape, gm.shape)
        report[f"v{i}.mask_agree"] = float((pm == gm).mean())
        gn = unpack_mask(g, f"v{i}.non_ambiguous_mask")
        report[f"v{i}.non_ambiguous_mask_agree"] = float((p["non_ambiguous_mask"].astype(bool) == gn).mean())
        both = (pm & gm)[:, ::stride, ::stride]
        for k in ("pts3d", "pts3d_cam", "ray_directions", "depth_along_ray", "depth_z", "conf",
                  "non_ambiguous_mask_logits", "img_no_norm"):
            a = p[k][:, ::stride, ::stride]
            b = g[f"v{i}.{k}"]
            assert a.shape == b.shape, (k, a.shape, b.shape)
            if k in MASKED_KEYS:
                a, b = a * both, b * both
            report[f"v{i}.{k}"] = rel_l2(a, b)
        for k in ("cam_trans", "camera_poses", "intrinsics", "metric_scaling_factor"):
            assert p[k].shape == g[f"v{i}.{k}"].shape, k
            report[f"v{i}.{k}"] = rel_l2(p[k], g[f"v{i}.{k}"])
        report[f"v{i}.cam_quats_deg"] = quat_angle_deg(p["cam_quats"], g[f"v{i}.cam_quats"])
    if tol_dense is None:
        return report
    bad = out_of_tolerance(report, tol_dense, tol_pose_deg, tol_small, min_mask_agree)
    assert not bad, f"out of tolerance: {bad}\nfull report: {report}"
    return report


SMALL_REPORT_KEYS = ("cam_trans", "camera_poses", "intrinsics", "metric_scaling_factor")
# quantities that contain the predicted camera pose (world-frame points, pose matrices): their error
# is dominated by the pose head's amplification of bf16 trunk noise, see DESIGN.md "Parity"
# "intrinsics" is a least-squares fit through ~50x50 ray samples (fewer on small images): it amplifies
# ray noise the same way, in both paths
POSE_DEPENDENT_KEYS = ("pts3d", "cam_trans", "camera_poses", "cam_quats_deg", "intrinsics")


def out_of_tolerance(report, tol_dense, tol_pose_deg, tol_small=None, min_mask_agree=0.0, yardstick=None,
                     slack=1.5):
    """Keys of `report` that exceed their tolerance.  With `yardstick` (the report of the oracle run
    with the reference's own bf16-autocast precision regions against the same fp32 golden), a key
    passes when it is within max(tolerance, slack * yardstick error): the CUDA path has to be as
    close to the fp32 reference as the reference's own bf16 path is."""
    bad = []
    # the two paths carry independent bf16 noise realisations, so the yardstick of a key is its
    # worst view in the scene, not the same view
    worst = {}
    for k, v in (yardstick or {}).items():
        name = k.split(".", 1)[1]
        worst[name] = min(worst.get(name, 1.0), v) if k.endswith("_agree") else max(worst.get(name, 0.0), v)
    for k, v in report.items():
        name = k.split(".", 1)[1]
        y = slack * worst[name] if name in worst else 0.0
        if k.endswith("_agree"):
            floor = min_mask_agree
            if name in worst:
                floor = min(floor, 1.0 - slack * (1.0 - worst[name]))
            if v < floor:
                bad.append((k, v))
        elif k.endswith("_deg"):
            if v > max(tol_pose_deg, y):
                bad.append((k, v))
        elif name in SMALL_REPORT_KEYS:
            if v > max(tol_small if tol_small is not None else tol_dense, y):
                bad.append((k, v))
        elif v > max(tol_dense, y):
            bad.append((k, v))
    return bad


class strict_fp32:
    """Context manager: no TF32 in cuBLAS / cuDNN, so that an oracle run on the GPU is true fp32
    (PyTorch allows TF32 cuDNN convolutions by default, which is ~1e-3 noise of its own)."""

    def __enter__(self):
        self.prev = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
        torch.backends.cuda.matmul.allow_tf32 = False
        torch.backends.cudnn.allow_tf32 = False
        return self

    def __exit__(self, *exc):
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = self.prev
        return False
