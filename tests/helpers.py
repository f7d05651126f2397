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
# end-to-end floor on the agreement of the boolean masks with the fp32 reference's, and how far beyond the
# reference's own bf16 path (worst view of the scene) a key may sit when it misses its absolute tolerance
MIN_MASK_AGREE = 0.985
YARDSTICK_SLACK = 1.5
# Pose-dependent keys (POSE_DEPENDENT_KEYS: world-frame points, pose matrices, translations, the pose angle)
# are a handful of numbers per view that the pose head amplifies out of ONE bf16 noise draw per path; the
# draw changes with any change of rounding anywhere upstream (fusing one multiply-add moved v0.pts3d of
# small_escaling from 0.5e-2 to 1.2e-2 while the yardstick's draw stayed at 0.4e-2), so a ratio of two single
# draws has a heavy tail: over the cases of profiles/r2_parity_report.json the CUDA path sits between 0.35x and
# 3.3x of the reference's own bf16 path (median 0.8x), and that path itself is at 0.4-2.7e-2 / 0.2-1.1 deg.
# They pass inside max(POSE_TOL, POSE_SLACK x yardstick), the angle inside max(0.1 deg, POSE_SLACK x
# yardstick) and never beyond MAX_POSE_DEG, so a pose-head regression (errors of order 1) cannot hide; the
# head's own arithmetic is held to the absolute BASELINE tolerances (0.1 deg, 1e-2) on identical inputs by
# test_heads_on_identical_inputs.
POSE_SLACK = 3.0
POSE_TOL = 1.5e-2
MAX_POSE_DEG = 2.0

def tol_intrinsics_for(H):
    """The intrinsics are a least-squares fit through ray samples on a stride-max(1, H // 50) grid
    (geometry.py:304-447): at 392x518 that is 56 x 52 samples and the fit is inside 1e-2 (measured 1-3e-3);
    on the 56x84 / 70x98 test images it amplifies the 2e-3 ray noise of EITHER path to 1-3e-2."""
    return None if H >= 392 else 3e-2


def out_of_tolerance(report, tol_dense, tol_pose_deg, tol_small=None, min_mask_agree=0.0, yardstick=None,
                     slack=YARDSTICK_SLACK, pose_slack=POSE_SLACK, tol_intrinsics=None):
    """Keys of `report` that exceed their tolerance.  With `yardstick` (the report of the oracle run
    with the reference's own bf16-autocast precision regions against the same fp32 golden), a key
    passes when it is within max(tolerance, slack * yardstick error) -- pose_slack for POSE_DEPENDENT_KEYS --:
    the CUDA path has to be as close to the fp32 reference as the reference's own bf16 path is."""
    bad = []
    # the two paths carry independent bf16 noise realisations, so the yardstick of a key is its
    # worst view in the scene, not the same view
    worst = {}
    for k, v in (yardstick or {}).items():
        name = k.split(".", 1)[1]
        worst[name] = min(worst.get(name, 1.0), v) if k.endswith("_agree") else max(worst.get(name, 0.0), v)
    for k, v in report.items():
        name = k.split(".", 1)[1]
        y = (pose_slack if name in POSE_DEPENDENT_KEYS else slack) * worst[name] if name in worst else 0.0
        if k.endswith("_deg") and v > MAX_POSE_DEG:
            bad.append((k, v))
            continue
        if k.endswith("_agree"):
            floor = min_mask_agree
            if name in worst:
                floor = min(floor, 1.0 - slack * (1.0 - worst[name]))
            if v < floor:
                bad.append((k, v))
        elif k.endswith("_deg"):
            if v > max(tol_pose_deg, y):
                bad.append((k, v))
        elif name == "intrinsics" and tol_intrinsics is not None:
            if v > max(tol_intrinsics, y):
                bad.append((k, v))
        elif name in POSE_DEPENDENT_KEYS:
            if v > max(POSE_TOL, y):
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


def parity_report_path():
    """Per-key errors the GPU parity tests measure are written here (JSON, one entry per test case) so
    that the numbers outlive the run: gpurun_out/ travels back from the GPU box, the copy that is
    judged is committed as profiles/r2_parity_report.json."""
    p = os.environ.get("MAB_PARITY_REPORT")
    if p:
        return p
    d = os.path.join(ROOT, "gpurun_out")
    os.makedirs(d, exist_ok=True)
    return os.path.join(d, "parity_report.json")


def record_parity(case, cuda_vs_fp32, yardstick_vs_fp32=None, extra=None):
    """Append / replace one case of the parity report: worst view per key of the CUDA path against the
    fp32 reference result, and of the reference-precision (bf16 autocast trunk, TF32 heads) oracle run
    against the same fp32 result."""
    import json

    def worst(rep):
        out = {}
        for k, v in (rep or {}).items():
            name = k.split(".", 1)[1] if "." in k and k[0] == "v" and k[1].isdigit() else k
            out[name] = min(out.get(name, 1.0), v) if name.endswith("_agree") else max(out.get(name, 0.0), v)
        return {k: float(v) for k, v in sorted(out.items())}

    path = parity_report_path()
    try:
        with open(path) as f:
            doc = json.load(f)
    except Exception:
        doc = {"what": "worst view per key: relative L2 (dense and small float keys), degrees (cam_quats_deg), "
                       "fraction of equal pixels (*_agree); 'cuda' = libmab200 path vs the fp32 reference result "
                       "(golden of the live reference, or the fp32 oracle on the same GPU), 'ref_bf16' = the "
                       "oracle with the reference's own precision regions (bf16 autocast trunk, TF32 heads) vs "
                       "the same fp32 result", "cases": {}}
    entry = {"cuda": worst(cuda_vs_fp32)}
    if yardstick_vs_fp32 is not None:
        entry["ref_bf16"] = worst(yardstick_vs_fp32)
    if extra:
        entry.update(extra)
    doc["cases"][case] = entry
    with open(path, "w") as f:
        json.dump(doc, f, indent=1, sort_keys=True)
    return entry


def compare_pred_lists(preds, ref, stride=1):
    """Report dict (same keys as compare_with_golden) of a prediction list against a reference
    prediction list (oracle output, torch tensors on any device)."""
    def npy(t):
        return t.detach().float().cpu().numpy() if t.dtype != torch.bool else t.detach().cpu().numpy()

    report = {}
    for i, (p, r) in enumerate(zip(preds, ref)):
        pm, rm = npy(p["mask"]).astype(bool), npy(r["mask"]).astype(bool)
        report[f"v{i}.mask_agree"] = float((pm == rm).mean())
        report[f"v{i}.non_ambiguous_mask_agree"] = float(
            (npy(p["non_ambiguous_mask"]).astype(bool) == npy(r["non_ambiguous_mask"]).astype(bool)).mean())
        both = (pm & rm)[:, ::stride, ::stride]
        for k in ("pts3d", "pts3d_cam", "ray_directions", "depth_along_ray", "depth_z", "conf",
                  "non_ambiguous_mask_logits", "img_no_norm"):
            a, b = npy(p[k])[:, ::stride, ::stride], npy(r[k])[:, ::stride, ::stride]
            assert a.shape == b.shape, (k, a.shape, b.shape)
            if k in MASKED_KEYS:
                a, b = a * both.reshape(both.shape[:3] + (1,) * (a.ndim - 3)), b * both.reshape(both.shape[:3] + (1,) * (b.ndim - 3))
            report[f"v{i}.{k}"] = rel_l2(a, b)
        for k in SMALL_REPORT_KEYS:
            report[f"v{i}.{k}"] = rel_l2(npy(p[k]), npy(r[k]))
        report[f"v{i}.cam_quats_deg"] = quat_angle_deg(npy(p["cam_quats"]), npy(r["cam_quats"]))
    return report
