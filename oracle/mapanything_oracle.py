"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's MapAnything inference path.

A functional, fp32, plain-PyTorch restatement of exactly what
`MapAnything.infer(views)` (/root/reference/mapanything/models/mapanything/model.py:2206-2355)
computes for the production configuration.  Every function cites the
reference file:line it follows.  It consumes a state dict in the reference key space
(oracle/state_dict_manifest.json) and is PINNED against the live reference by
oracle/make_golden.py -> tests/golden/*.npz (generated in the build container, where
/root/reference is importable) and tests/test_oracle_golden.py.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product (mapanything_b200) never does; it fails loudly without its CUDA
library.
"""
import math
from typing import Any, Dict, List, Optional

import numpy as np
import torch
import torch.nn.functional as F

PATCH = 14
LN_EPS = 1e-6  # partial(nn.LayerNorm, eps=1e-6): vision_transformer.py:110, aat.py:44
IMG_MEAN = (0.485, 0.456, 0.406)  # uniception/models/encoders/image_normalizations.py:21
IMG_STD = (0.229, 0.224, 0.225)


def _ln(x, sd, prefix, eps=LN_EPS):
    return F.layer_norm(x, (x.shape[-1],), sd[prefix + ".weight"], sd[prefix + ".bias"], eps)


def _lin(x, sd, prefix):
    return F.linear(x, sd[prefix + ".weight"], sd.get(prefix + ".bias"))


def _conv(x, sd, prefix, stride=1, padding=0):
    return F.conv2d(x, sd[prefix + ".weight"], sd.get(prefix + ".bias"), stride=stride, padding=padding)


def _attention(x, sd, prefix, num_heads, entropy_scaling=False):
    """transformer_blocks.py:174-212 / encoders/dinov2.py:128-141 (SDPA, scale = head_dim**-0.5).
    entropy_scaling: q *= sqrt(1.4 * log(N) / log(444)) (transformer_blocks.py:108-110,190-196)."""
    B, N, C = x.shape
    hd = C // num_heads
    qkv = _lin(x, sd, prefix + ".qkv").reshape(B, N, 3, num_heads, hd).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0], qkv[1], qkv[2]
    if entropy_scaling:
        q = q * torch.sqrt((1.4 * torch.log(torch.tensor(N, device=q.device))) /
                           torch.log(torch.tensor(444, device=q.device)))
    o = F.scaled_dot_product_attention(q, k, v, scale=hd ** -0.5)
    o = o.transpose(1, 2).reshape(B, N, C)
    return _lin(o, sd, prefix + ".proj")


def _mlp(x, sd, prefix):
    """transformer_blocks.py:81-88 / dinov2 layers/mlp.py:34-40 (exact-erf GELU)."""
    return _lin(F.gelu(_lin(x, sd, prefix + ".fc1")), sd, prefix + ".fc2")


# --------------------------------------------------------------------------------------------
# a3-a5: DINOv2 ViT-L/14 encoder
# --------------------------------------------------------------------------------------------
def interpolate_pos_embed(pos_embed, h0, w0, offset=0.1):
    """vision_transformer.py:208-242 (`w`,`h` there are H,W of the image; interpolate_offset=0.1,
    antialias False, bicubic with explicit scale factors)."""
    N = pos_embed.shape[1] - 1
    M = int(math.sqrt(N))
    assert M * M == N
    dim = pos_embed.shape[-1]
    cls_pos = pos_embed[:, 0]
    patch_pos = pos_embed[:, 1:]
    if h0 * w0 == N and h0 == w0:
        return pos_embed
    sx = float(h0 + offset) / M
    sy = float(w0 + offset) / M
    patch_pos = F.interpolate(patch_pos.reshape(1, M, M, dim).permute(0, 3, 1, 2), mode="bicubic",
                              antialias=False, scale_factor=(sx, sy))
    assert patch_pos.shape[-2:] == (h0, w0)
    patch_pos = patch_pos.permute(0, 2, 3, 1).reshape(1, -1, dim)
    return torch.cat((cls_pos.unsqueeze(0), patch_pos), dim=1)


def dinov2_encoder(imgs, sd, prefix="encoder.model", taps=None):
    """encoders/dinov2.py:146-178 -> vision_transformer.py:244-312, block.py:93-118.
    imgs (V,3,H,W) -> patch tokens (V, h*w, 1024), token-major (the reference then permutes to
    NCHW, dinov2.py:173-176; the oracle keeps token-major and converts at the API edge)."""
    V, _, H, W = imgs.shape
    assert H % PATCH == 0 and W % PATCH == 0
    h0, w0 = H // PATCH, W // PATCH
    x = F.conv2d(imgs, sd[prefix + ".patch_embed.proj.weight"], sd[prefix + ".patch_embed.proj.bias"],
                 stride=PATCH)  # patch_embed.py:80
    x = x.flatten(2).transpose(1, 2)
    x = torch.cat((sd[prefix + ".cls_token"].expand(V, -1, -1), x), dim=1)
    x = x + interpolate_pos_embed(sd[prefix + ".pos_embed"], h0, w0)
    if taps is not None:
        taps["enc_tokens_in"] = x
    depth = 1 + max(int(k.split(".")[3]) for k in sd if k.startswith(prefix + ".blocks."))
    for i in range(depth):
        p = f"{prefix}.blocks.{i}"
        x = x + sd[p + ".ls1.gamma"] * _attention(_ln(x, sd, p + ".norm1"), sd, p + ".attn", 16)
        x = x + sd[p + ".ls2.gamma"] * _mlp(_ln(x, sd, p + ".norm2"), sd, p + ".mlp")
        if taps is not None and i == 0:
            taps["enc_block0"] = x
    x = _ln(x, sd, prefix + ".norm")
    return x[:, 1:]  # x_norm_patchtokens


# --------------------------------------------------------------------------------------------
# a1: pre-processing  (mapanything/utils/inference.py:222-311, geometry.py)
# --------------------------------------------------------------------------------------------
def get_rays_in_camera_frame(K, H, W):
    """geometry.py:186-241, normalize_to_unit_sphere=True. K (B,3,3) -> (B,H,W,3)."""
    xs = torch.arange(W, dtype=torch.float32, device=K.device)
    ys = torch.arange(H, dtype=torch.float32, device=K.device)
    xg, yg = torch.meshgrid(xs, ys, indexing="xy")
    fx, fy = K[:, 0, 0].view(-1, 1, 1), K[:, 1, 1].view(-1, 1, 1)
    cx, cy = K[:, 0, 2].view(-1, 1, 1), K[:, 1, 2].view(-1, 1, 1)
    xx = (xg.unsqueeze(0) - cx) / fx
    yy = (yg.unsqueeze(0) - cy) / fy
    r = torch.stack((xx, yy, torch.ones_like(xx)), dim=-1)
    return r / torch.norm(r, dim=-1, keepdim=True)


def quaternion_to_rotation_matrix(q):
    """geometry.py:601-652 (xyzw, re-normalises)."""
    q = q / q.norm(dim=1, keepdim=True)
    x, y, z, w = q.unbind(dim=1)
    return torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], dim=1).view(-1, 3, 3)


def rotation_matrix_to_quaternion(m):
    """geometry.py:655-713 (pytorch3d-style, output xyzw with w >= 0)."""
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = torch.unbind(m.reshape(m.shape[:-2] + (9,)), dim=-1)
    x = torch.stack([1.0 + m00 + m11 + m22, 1.0 + m00 - m11 - m22,
                     1.0 - m00 + m11 - m22, 1.0 - m00 - m11 + m22], dim=-1)
    q_abs = torch.where(x > 0, torch.sqrt(x.clamp(min=0)), torch.zeros_like(x))
    cand = torch.stack([
        torch.stack([q_abs[..., 0] ** 2, m21 - m12, m02 - m20, m10 - m01], dim=-1),
        torch.stack([m21 - m12, q_abs[..., 1] ** 2, m10 + m01, m02 + m20], dim=-1),
        torch.stack([m02 - m20, m10 + m01, q_abs[..., 2] ** 2, m12 + m21], dim=-1),
        torch.stack([m10 - m01, m20 + m02, m21 + m12, q_abs[..., 3] ** 2], dim=-1)], dim=-2)
    cand = cand / (2.0 * q_abs[..., None].clamp(min=0.1))
    idx = q_abs.argmax(dim=-1)
    out = torch.gather(cand, -2, idx[..., None, None].expand(idx.shape + (1, 4))).squeeze(-2)
    out = out[..., [1, 2, 3, 0]]
    return torch.where(out[..., 3:4] < 0, -out, out)


def quaternion_multiply(q1, q2):
    """geometry.py:774-808."""
    x1, y1, z1, w1 = q1.unbind(dim=1)
    x2, y2, z2, w2 = q2.unbind(dim=1)
    return torch.stack([w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                        w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                        w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2,
                        w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2], dim=1)


def transform_pose_2_to_1(q1, t1, q2, t2):
    """geometry.py:814-852: pose2 expressed in the frame of pose1."""
    qc = q1.clone()
    qc[:, :3] = -qc[:, :3]
    inv_q1 = qc / torch.sum(q1 * q1, dim=1, keepdim=True)  # geometry.py:744-772
    R1_inv = quaternion_to_rotation_matrix(inv_q1)
    t1_inv = -torch.einsum("bij,bj->bi", R1_inv, t1)
    return quaternion_multiply(inv_q1, q2), torch.einsum("bij,bj->bi", R1_inv, t2) + t1_inv


def preprocess_views(views):
    """inference.py:222-311."""
    out = []
    for view in views:
        pv = dict(view)
        if "intrinsics" in view:
            H, W = view["img"].shape[-2:]
            pv["ray_directions_cam"] = get_rays_in_camera_frame(view["intrinsics"], H, W)
            del pv["intrinsics"]
        elif "ray_directions" in view:
            r = view["ray_directions"]
            pv["ray_directions_cam"] = r / (torch.norm(r, dim=-1, keepdim=True) + 1e-8)
            del pv["ray_directions"]
        if "depth_z" in view:
            r = pv["ray_directions_cam"]
            # NB: the reference indexes depth_z[..., None], i.e. depth_z is (B,H,W) in practice
            pts = view["depth_z"][..., None] * (r / r[..., 2:3])
            pv["depth_along_ray"] = torch.norm(pts, dim=-1, keepdim=True)
            del pv["depth_z"]
        if "camera_poses" in view:
            cp = view["camera_poses"]
            if isinstance(cp, tuple) and len(cp) == 2:
                pv["camera_pose_quats"], pv["camera_pose_trans"] = cp
            elif torch.is_tensor(cp) and cp.shape[-2:] == (4, 4):
                pv["camera_pose_quats"] = rotation_matrix_to_quaternion(cp[:, :3, :3])
                pv["camera_pose_trans"] = cp[:, :3, 3]
            else:
                raise ValueError("camera_poses must be either a tuple of (quats, trans) or (B,4,4)")
            del pv["camera_poses"]
        if "is_metric_scale" not in pv:
            pv["is_metric_scale"] = torch.ones(view["img"].shape[0], dtype=torch.bool)
        out.append(pv)
    return out


# --------------------------------------------------------------------------------------------
# a6: geometric-input encoders and fusion (model.py:792-1438)
# --------------------------------------------------------------------------------------------
def dense_rep_encoder(x, sd, prefix):
    """dense_rep_encoder.py:234-287: PixelUnshuffle(14) -> conv3x3 -> 2 ResidualBlocks -> conv1x1
    -> LayerNorm.  x (V,C,H,W) -> (V, h*w, 1024) token-major."""
    x = F.pixel_unshuffle(x, PATCH)
    x = _conv(x, sd, prefix + ".conv_in", padding=1)
    for i in (0, 1):  # ResidualBlock, dense_rep_encoder.py:31-52 (GELU)
        p = f"{prefix}.encoder.{i}"
        idt = _conv(x, sd, p + ".shortcut")
        o = _conv(F.gelu(_conv(x, sd, p + ".conv1", padding=1)), sd, p + ".conv2", padding=1)
        x = F.gelu(o + idt)
    x = _conv(x, sd, prefix + ".encoder.2")
    x = x.flatten(2).transpose(1, 2)
    return F.layer_norm(x, (x.shape[-1],), sd[prefix + ".norm_layer.weight"],
                        sd[prefix + ".norm_layer.bias"], LN_EPS)  # dense_rep_encoder.py:69


def global_rep_encoder(x, sd, prefix):
    """global_rep_encoder.py:54-67,83-105: Linear-GELU x3 -> Linear -> LayerNorm(eps 1e-6, :24)."""
    for name in ("encoder.0.0.0.0", "encoder.0.0.1", "encoder.0.1"):
        x = F.gelu(_lin(x, sd, f"{prefix}.{name}"))
    x = _lin(x, sd, f"{prefix}.encoder.1")
    return F.layer_norm(x, (x.shape[-1],), sd[prefix + ".norm_layer.weight"],
                        sd[prefix + ".norm_layer.bias"], LN_EPS)


def fuse_geometric_inputs(feats, views, sd, flags):
    """model.py:1377-1436.  feats (V, T, 1024) image-encoder tokens (B=1 per view).
    flags: dict(use_calibration, use_depth, use_pose, use_depth_scale, use_pose_scale) following
    model.py:2154-2197; the per-sample masks become deterministic 0/1 at inference."""
    V = len(views)
    H, W = views[0]["img"].shape[-2:]
    dev = feats.device
    any_geo = flags["use_calibration"] or flags["use_depth"] or flags["use_pose"]
    # model.py:1336-1369 (masks) + per-view availability checks :936-949, :1036-1117, :826-855
    ray_mask = torch.tensor([any_geo and flags["use_calibration"] and "ray_directions_cam" in v
                             for v in views], device=dev)
    depth_mask = torch.tensor([any_geo and flags["use_depth"] and "depth_along_ray" in v
                               for v in views], device=dev)
    cam_mask = torch.tensor([any_geo and flags["use_pose"] and "camera_pose_quats" in v
                             and "camera_pose_trans" in v for v in views], device=dev)

    # poses relative to view 0 (model.py:792-896)
    quats = torch.tensor([0.0, 0.0, 0.0, 1.0], device=dev).repeat(V, 1)
    trans = torch.zeros(V, 3, device=dev)
    if cam_mask.any():
        idx = cam_mask.nonzero().flatten().tolist()
        q2 = torch.cat([views[i]["camera_pose_quats"] for i in idx], 0).float()
        t2 = torch.cat([views[i]["camera_pose_trans"] for i in idx], 0).float()
        q1 = views[0]["camera_pose_quats"].float().repeat(len(idx), 1)
        t1 = views[0]["camera_pose_trans"].float().repeat(len(idx), 1)
        q, t = transform_pose_2_to_1(q1, t1, q2, t2)
        quats[cam_mask], trans[cam_mask] = q, t

    if ray_mask.any():  # model.py:898-970 (skipped when all-False: result is exactly 0)
        rays = torch.zeros(V, H, W, 3, device=dev)
        for i in range(V):
            if ray_mask[i]:
                rays[i] = views[i]["ray_directions_cam"][0]
        rf = dense_rep_encoder(rays.permute(0, 3, 1, 2).contiguous(), sd, "ray_dirs_encoder")
        feats = feats + rf * ray_mask.view(V, 1, 1)

    if depth_mask.any():  # model.py:973-1168
        depth = torch.zeros(V, H, W, 1, device=dev)
        nf = torch.zeros(V, device=dev)
        metric = torch.zeros(V, dtype=torch.bool, device=dev)
        for i in range(V):
            if depth_mask[i]:
                d = views[i]["depth_along_ray"].float()
                m = views[i].get("is_metric_scale", torch.zeros(1, dtype=torch.bool)).clone()
                if not flags["use_depth_scale"]:
                    m[:] = False  # depth_scale_norm_all_prob = 1 (model.py:1055-1060)
                valid = d > 0  # normalize_depth_using_non_zero_pixels geometry.py:1594-1626
                f_ = (d * valid).sum(dim=(1, 2, 3)) / (valid.sum(dim=(1, 2, 3)) + 1e-8)
                f_ = f_.clip(min=1e-8)
                depth[i] = (d / f_.view(-1, 1, 1, 1))[0]
                nf[i] = f_[0]
                metric[i] = bool(m[0])
        org = depth.norm(dim=-1, keepdim=True)  # apply_log_to_norm geometry.py:1737-1750
        depth = depth / org.clip(min=1e-8) * torch.log1p(org)
        df = dense_rep_encoder(depth.permute(0, 3, 1, 2).contiguous(), sd, "depth_encoder")
        df = df * depth_mask.view(V, 1, 1)
        sf = global_rep_encoder(torch.log(nf + 1e-8).unsqueeze(-1), sd, "depth_scale_encoder")
        sf = sf * depth_mask.view(V, 1) * metric.view(V, 1)
        feats = feats + df + sf.unsqueeze(1)

    if cam_mask.any():  # model.py:1170-1289
        qf = global_rep_encoder(quats, sd, "cam_rot_encoder") * cam_mask.view(V, 1)
        metric = torch.tensor([bool(v["is_metric_scale"][0]) if "is_metric_scale" in v else False
                               for v in views], device=dev)
        if not flags["use_pose_scale"]:
            metric[:] = False
        dis = trans.norm(dim=-1)  # normalize_pose_translations geometry.py:1629-1666 (over views)
        nfac = (dis.sum() / ((dis > 0).sum() + 1e-8)).clip(min=1e-8)
        tf = global_rep_encoder(trans / nfac, sd, "cam_trans_encoder") * cam_mask.view(V, 1)
        tsf = global_rep_encoder(torch.log(nfac + 1e-8).view(1, 1).repeat(V, 1), sd,
                                 "cam_trans_scale_encoder")
        tsf = tsf * cam_mask.view(V, 1) * metric.view(V, 1)
        feats = feats + (qf + tf + tsf).unsqueeze(1)
    else:
        # the reference always runs the three pose MLPs and multiplies by the (False) mask: adds 0
        pass
    return _ln(feats, sd, "fusion_norm_layer")  # model.py:106-108,1423-1431 LayerNorm(1024, eps 1e-6)


# --------------------------------------------------------------------------------------------
# a7-a9: multi-view alternating-attention transformer (IFR)
# --------------------------------------------------------------------------------------------
def _block(x, sd, p, heads, entropy_scaling=False):
    """transformer_blocks.py:467-468 (no LayerScale: init_values=None)."""
    x = x + _attention(_ln(x, sd, p + ".norm1"), sd, p + ".attn", heads, entropy_scaling)
    return x + _mlp(_ln(x, sd, p + ".norm2"), sd, p + ".mlp")


# info_sharing module_args that change the arithmetic (configs/model/info_sharing/*.yaml):
#   use_entropy_scaling (aat_ifr_24_layers_escaling.yaml), distinguish_ref_and_non_ref_views
#   (aat_ifr_24_layers_no_ref_view.yaml).  Set by infer(..., aat_opts=...).
#   use_pe_for_non_reference_views (aat.py:30,628-645): every non-reference view adds a row of the
#   view_pos_table; the rows are drawn with torch.randint(1, max_num_views_for_pe) from the global
#   generator at every call (use_rand_idx_pe_for_non_reference_views, the default) or are 1..V-1.
AAT_OPTS = {"use_entropy_scaling": False, "distinguish_ref_and_non_ref_views": True, "indices": (11, 17),
            "use_pe_for_non_reference_views": False, "max_num_views_for_pe": 1000,
            "use_rand_idx_pe_for_non_reference_views": True,
            # model_type "global_attention" (global_attention_transformer.py:213-345,530-640): every block
            # attends over the whole scene and every view takes a row of the [max_num_views, dim] table
            "all_global": False}


def aat_opts_for(info_sharing_config):
    """AAT_OPTS overrides of an info_sharing config dict (model_type + module_args)."""
    margs = dict(info_sharing_config.get("module_args", {}))
    if info_sharing_config.get("model_type", "alternating_attention") == "global_attention":
        margs.update(all_global=True, distinguish_ref_and_non_ref_views=True, use_pe_for_non_reference_views=True,
                     max_num_views_for_pe=margs["max_num_views"])
    return margs


def alternating_attention_transformer(feats, sd, indices=None, prefix="info_sharing", taps=None):
    """alternating_attention_transformer.py:530-771.  feats (V, T, 1024) fused tokens of one scene
    (B=1).  Returns (final (V,T,dim), [tap (V,T,dim)...], scale_token_final (dim,)).  Width, head
    count (dim / 64: 768 / 12 and 1024 / 16 in the shipped configs) and depth come from the state dict."""
    V, T, _ = feats.shape
    indices = AAT_OPTS["indices"] if indices is None else indices
    x = torch.cat([feats.reshape(1, V * T, -1), sd["scale_token"].view(1, 1, -1)], dim=1)
    if prefix + ".proj_embed.weight" in sd:  # aat.py:121-124: nn.Identity when the widths agree
        x = _lin(x, sd, prefix + ".proj_embed")
    heads = sd[prefix + ".norm.weight"].numel() // 64
    if AAT_OPTS["distinguish_ref_and_non_ref_views"]:  # aat.py:618-626
        x = torch.cat([x[:, :T] + sd[prefix + ".view_pos_table"][0].view(1, 1, -1), x[:, T:]], dim=1)
        if AAT_OPTS["use_pe_for_non_reference_views"]:  # aat.py:554-557,628-645
            assert V <= AAT_OPTS["max_num_views_for_pe"]
            if AAT_OPTS["use_rand_idx_pe_for_non_reference_views"]:
                rows = torch.randint(low=1, high=AAT_OPTS["max_num_views_for_pe"], size=(V - 1,))
            else:
                rows = torch.arange(1, V)
            pe = sd[prefix + ".view_pos_table"][rows.to(x.device)].repeat_interleave(T, dim=0).unsqueeze(0)
            x = torch.cat([x[:, :T], x[:, T:V * T] + pe, x[:, V * T:]], dim=1)
    es = AAT_OPTS["use_entropy_scaling"]
    if taps is not None:
        taps["aat_in"] = x[0]
    depth = 1 + max(int(k.split(".")[2]) for k in sd if k.startswith(prefix + ".self_attention_blocks."))
    inter = []
    for i in range(depth):
        p = f"{prefix}.self_attention_blocks.{i}"
        if AAT_OPTS["all_global"] or i % 2 == 0:
            x = _block(x, sd, p, heads, es)
        else:
            extra = x[:, V * T:]
            y = _block(x[:, :V * T].reshape(V, T, -1), sd, p, heads, es)
            x = torch.cat([y.reshape(1, V * T, -1), extra], dim=1)
        if taps is not None and i in (0, 1):
            taps[f"aat_block{i}"] = x[0]
        if i in indices:
            inter.append(_ln(x, sd, prefix + ".norm"))
    out = _ln(x, sd, prefix + ".norm")
    dim = out.shape[-1]
    return (out[0, :V * T].reshape(V, T, dim), [t[0, :V * T].reshape(V, T, dim) for t in inter],
            out[0, V * T])


# --------------------------------------------------------------------------------------------
# a10-a14: heads
# --------------------------------------------------------------------------------------------
def _rcu(x, sd, p):
    """dpt_block.py:156-177 ResidualConvUnit_custom (ReLU, no bn)."""
    o = _conv(F.relu(x), sd, p + ".conv1", padding=1)
    o = _conv(F.relu(o), sd, p + ".conv2", padding=1)
    return o + x


def _fusion(sd, p, x, skip=None):
    """dpt_block.py:225-255 FeatureFusionBlock_custom (align_corners=True, width_ratio=1)."""
    if skip is not None:
        x = x + _rcu(skip, sd, p + ".resConfUnit1")
    x = _rcu(x, sd, p + ".resConfUnit2")
    x = F.interpolate(x, scale_factor=2, mode="bilinear", align_corners=True)
    return _conv(x, sd, p + ".out_conv")


def dpt_feature(layers, sd, prefix="dpt_feature_head"):
    """dpt.py:180-232.  layers: 4 NCHW tensors [(V,1024,h,w),(V,768,h,w)x3] -> (V,256,8h,8w)."""
    ip = prefix + ".input_process"
    l0 = F.conv_transpose2d(_conv(layers[0], sd, ip + ".0.0.0"), sd[ip + ".0.0.1.weight"],
                            sd[ip + ".0.0.1.bias"], stride=4)
    l1 = F.conv_transpose2d(_conv(layers[1], sd, ip + ".1.0.0"), sd[ip + ".1.0.1.weight"],
                            sd[ip + ".1.0.1.bias"], stride=2)
    l2 = _conv(layers[2], sd, ip + ".2.0.0")
    l3 = _conv(_conv(layers[3], sd, ip + ".3.0.0"), sd, ip + ".3.0.1", stride=2, padding=1)
    sc = prefix + ".scratch"
    l0 = F.conv2d(l0, sd[sc + ".layer1_rn.weight"], padding=1)
    l1 = F.conv2d(l1, sd[sc + ".layer2_rn.weight"], padding=1)
    l2 = F.conv2d(l2, sd[sc + ".layer3_rn.weight"], padding=1)
    l3 = F.conv2d(l3, sd[sc + ".layer4_rn.weight"], padding=1)
    p4 = _fusion(sd, sc + ".refinenet4", l3)[:, :, :l2.shape[2], :l2.shape[3]]
    p3 = _fusion(sd, sc + ".refinenet3", p4, l2)
    p2 = _fusion(sd, sc + ".refinenet2", p3, l1)
    return _fusion(sd, sc + ".refinenet1", p2, l0)


def dpt_regressor(x, sd, out_hw, prefix="dpt_regressor_head"):
    """dpt.py:285-311."""
    x = _conv(x, sd, prefix + ".conv1", padding=1)
    x = F.interpolate(x, size=out_hw, mode="bilinear", align_corners=True)
    x = F.relu(_conv(x, sd, prefix + ".conv2.0", padding=1))
    return _conv(x, sd, prefix + ".conv2.2")


def pose_head(feat, sd, prefix="pose_head"):
    """pose_head.py:115-159 + adaptors.py:714-732 (linear trans, unit quats). feat (V,768,h,w)."""
    x = _conv(feat, sd, prefix + ".proj")
    for i in (0, 1):
        p = f"{prefix}.res_conv.{i}"
        y = F.relu(_conv(x, sd, p + ".res_conv1"))
        y = F.relu(_conv(y, sd, p + ".res_conv2"))
        y = F.relu(_conv(y, sd, p + ".res_conv3"))
        x = x + y
    x = x.mean(dim=(2, 3))
    x = F.relu(_lin(x, sd, prefix + ".more_mlps.0"))
    x = F.relu(_lin(x, sd, prefix + ".more_mlps.2"))
    t = _lin(x, sd, prefix + ".fc_t")
    q = _lin(x, sd, prefix + ".fc_rot")
    q = q / q.norm(dim=1, keepdim=True).clip(min=1e-8)
    return t, q


def scale_head(tok, sd, prefix="scale_head"):
    """mlp_head.py:58-92 + ScaleAdaptor adaptors.py:190-212 (exp, clip >= 1e-8). tok (768,)."""
    x = _lin(tok.view(1, -1), sd, prefix + ".proj")
    x = F.relu(_lin(x, sd, prefix + ".mlp.0.0"))
    x = F.relu(_lin(x, sd, prefix + ".mlp.1.0"))
    x = _lin(x, sd, prefix + ".output_proj")
    return torch.exp(x).clip(min=1e-8)  # (1,1)


# --------------------------------------------------------------------------------------------
# forward: model.py:1657-2152
# --------------------------------------------------------------------------------------------
DEFAULT_FLAGS = dict(use_calibration=True, use_depth=True, use_pose=True, use_depth_scale=True,
                     use_pose_scale=True)


import contextlib


def _trunk_ctx(amp, device_type):
    """amp=True reproduces the precision regions of the reference's default GPU path: encoder and
    info-sharing under torch.autocast(bf16) (model.py:2335-2338), everything else with autocast
    disabled (model.py:1377, 1774)."""
    if amp:
        return torch.autocast(device_type, dtype=torch.bfloat16)
    return contextlib.nullcontext()


def forward(sd, views, flags=None, taps=None, amp=False):
    """See _forward_impl.  amp=True on CUDA also allows TF32 in the fp32 regions, as the reference
    does at import time (model.py:89-93; cuDNN conv TF32 is PyTorch's default)."""
    if not (amp and views[0]["img"].device.type == "cuda"):
        return _forward_impl(sd, views, flags, taps, amp)
    prev = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.allow_tf32 = True
    try:
        return _forward_impl(sd, views, flags, taps, amp)
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = prev


def _forward_impl(sd, views, flags=None, taps=None, amp=False):
    """Raw (un-post-processed) per-view prediction dicts; views already pre-processed; B=1/view.
    amp=False: fp32 everywhere (infer(use_amp=False)); amp=True: the reference's bf16-autocast
    regions (used on the GPU box as the yardstick for "the reference's own bf16 path")."""
    flags = dict(DEFAULT_FLAGS, **(flags or {}))
    V = len(views)
    assert all(v["img"].shape == views[0]["img"].shape for v in views)
    B, _, H, W = views[0]["img"].shape
    assert B == 1, "the oracle restates the B=1-per-view case used by every config"
    h, w = H // PATCH, W // PATCH
    imgs = torch.cat([v["img"] for v in views], dim=0).float()
    dev_type = imgs.device.type
    with _trunk_ctx(amp, dev_type):
        enc = dinov2_encoder(imgs, sd, taps=taps).float()  # (V,T,1024)
    if taps is not None:
        taps["enc_out"] = enc
    fused = fuse_geometric_inputs(enc, views, sd, flags)
    if taps is not None:
        taps["fused"] = fused
    with _trunk_ctx(amp, dev_type):
        final, inter, scale_tok = alternating_attention_transformer(fused, sd, taps=taps)
    final, inter, scale_tok = final.float(), [t.float() for t in inter], scale_tok.float()
    if taps is not None:
        taps["aat_tap11"], taps["aat_tap17"], taps["aat_final"] = inter[0], inter[1], final
        taps["aat_inter"] = list(inter)
        taps["scale_token_final"] = scale_tok

    def nchw(t):
        return t.reshape(V, h, w, -1).permute(0, 3, 1, 2).contiguous()

    if len(inter) == 2:
        layers = [nchw(fused), nchw(inter[0]), nchw(inter[1]), nchw(final)]  # model.py:1724-1747
    else:  # three intermediate features: the DPT head does not see the encoder output (model.py:324-327,1748-1768)
        layers = [nchw(inter[0]), nchw(inter[1]), nchw(inter[2]), nchw(final)]
    feat8 = dpt_feature(layers, sd)
    logits = dpt_regressor(feat8, sd, (H, W))  # (V,6,H,W)
    cam_t, cam_q = pose_head(layers[3], sd)
    scale = scale_head(scale_tok, sd)  # (1,1)
    if taps is not None:
        taps["dpt_feat8"], taps["dense_logits"] = feat8, logits
        taps["pose_t"], taps["pose_q"], taps["scale"] = cam_t, cam_q, scale

    # dense adaptor: adaptors.py:1772-1797 (unit-sphere rays, exp depth, 1+exp conf, sigmoid mask)
    rays = logits[:, 0:3]
    rays = rays / rays.norm(dim=1, keepdim=True).clip(min=1e-8)
    depth = torch.exp(logits[:, 3:4]).clip(min=0)
    conf = 1 + torch.exp(logits[:, 4])
    mask_logits = logits[:, 5]
    mask = torch.sigmoid(mask_logits) > 0.5

    rays = rays.permute(0, 2, 3, 1).contiguous()
    depth = depth.permute(0, 2, 3, 1).contiguous()
    # geometry.py:855-907
    R = quaternion_to_rotation_matrix(cam_q / torch.norm(cam_q, dim=-1, keepdim=True))
    pts_cam = depth * rays
    pts = torch.einsum("bik,bhwk->bhwi", R, pts_cam) + cam_t.view(V, 1, 1, 3)
    res = []
    for i in range(V):  # model.py:1907-1923, 2117-2150
        res.append({
            "pts3d": pts[i:i + 1] * scale.view(1, 1, 1, 1),
            "pts3d_cam": pts_cam[i:i + 1] * scale.view(1, 1, 1, 1),
            "ray_directions": rays[i:i + 1],
            "depth_along_ray": depth[i:i + 1] * scale.view(1, 1, 1, 1),
            "cam_trans": cam_t[i:i + 1] * scale,
            "cam_quats": cam_q[i:i + 1],
            "metric_scaling_factor": scale,
            "conf": conf[i:i + 1],
            "non_ambiguous_mask": mask[i:i + 1],
            "non_ambiguous_mask_logits": mask_logits[i:i + 1],
        })
    return res


# --------------------------------------------------------------------------------------------
# a16: post-processing  (inference.py:314-506)
# --------------------------------------------------------------------------------------------
def recover_pinhole_intrinsics(rays):
    """geometry.py:304-447, regression branch (H*W <= 1e6). rays (B,H,W,3) -> (B,3,3)."""
    B, H, W, _ = rays.shape
    assert H * W <= 1000000
    hi = torch.arange(0, H, max(1, H // 50))
    wi = torch.arange(0, W, max(1, W // 50))
    rs = rays[:, hi[:, None], wi[None, :], :]
    xg = wi.float()[None, :].expand(len(hi), -1).reshape(1, -1).expand(B, -1)
    yg = hi.float()[:, None].expand(-1, len(wi)).reshape(1, -1).expand(B, -1)
    rx = (rs[..., 0] / rs[..., 2]).reshape(B, -1)
    ry = (rs[..., 1] / rs[..., 2]).reshape(B, -1)

    def solve(ratio, target):
        A = torch.stack([torch.ones_like(ratio), ratio], dim=2)
        ATA = torch.bmm(A.transpose(1, 2), A)
        ATb = torch.bmm(A.transpose(1, 2), target.unsqueeze(2))
        s = torch.linalg.solve(ATA, ATb).squeeze(2)
        return s[:, 0], s[:, 1]

    cx, fx = solve(rx, xg)
    cy, fy = solve(ry, yg)
    K = torch.zeros(B, 3, 3)
    K[:, 0, 0], K[:, 1, 1], K[:, 0, 2], K[:, 1, 2], K[:, 2, 2] = fx, fy, cx, cy, 1.0
    return K


def points_to_normals(point, mask):
    """geometry.py:1788-1851 (numpy, with mask, no edge_threshold)."""
    H, W = point.shape[:2]
    mp = np.zeros((H + 2, W + 2), dtype=bool)
    mp[1:-1, 1:-1] = mask
    pts = np.zeros((H + 2, W + 2, 3), dtype=point.dtype)
    pts[1:-1, 1:-1] = point
    c = pts[1:-1, 1:-1]
    up, left = pts[:-2, 1:-1] - c, pts[1:-1, :-2] - c
    down, right = pts[2:, 1:-1] - c, pts[1:-1, 2:] - c
    with np.errstate(all="ignore"):
        n = np.stack([np.cross(up, left), np.cross(left, down), np.cross(down, right), np.cross(right, up)])
        n = n / (np.linalg.norm(n, axis=-1, keepdims=True) + 1e-12)
        valid = np.stack([mp[:-2, 1:-1] & mp[1:-1, :-2], mp[1:-1, :-2] & mp[2:, 1:-1],
                          mp[2:, 1:-1] & mp[1:-1, 2:], mp[1:-1, 2:] & mp[:-2, 1:-1]]) & mp[None, 1:-1, 1:-1]
        n = (n * valid[..., None]).sum(axis=0)
        n = n / (np.linalg.norm(n, axis=-1, keepdims=True) + 1e-12)
    nm = valid.any(axis=0)
    return np.where(nm[..., None], n, 0), nm


def _win3(x, pad_mode, fill=None):
    """3x3 sliding windows (stride 1) -> (H,W,...,9) stacked on a new last axis."""
    if pad_mode == "edge":
        xp = np.pad(x, [(1, 1), (1, 1)] + [(0, 0)] * (x.ndim - 2), mode="edge")
    else:
        xp = np.pad(x, [(1, 1), (1, 1)] + [(0, 0)] * (x.ndim - 2), mode="constant", constant_values=fill)
    H, W = x.shape[:2]
    return np.stack([xp[dy:dy + H, dx:dx + W] for dy in range(3) for dx in range(3)], axis=-1)


def _maxpool3(x):
    """max_pool_2d(kernel 3, stride 1, padding 1) with NaN padding + nanmax, geometry.py:1965-2099."""
    with np.errstate(all="ignore"):
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            # separable like the reference (axis -2 then -1); nanmax of all-NaN -> NaN
            w = _win3(x, "constant", np.nan)
            return np.nanmax(w, axis=-1)


def depth_edge(depth, rtol, mask):
    """geometry.py:2102-2143."""
    with np.errstate(all="ignore"):
        diff = _maxpool3(np.where(mask, depth, -np.inf)) + _maxpool3(np.where(mask, -depth, -np.inf))
        return (diff / depth) > rtol


def normals_edge(normals, tol, mask):
    """geometry.py:2200-2258."""
    with np.errstate(all="ignore"):
        normals = normals / (np.linalg.norm(normals, axis=-1, keepdims=True) + 1e-12)
        nw = _win3(normals, "edge")  # (H,W,3,9)
        mw = _win3(mask, "edge")  # (H,W,9)
        ang = np.where(mw, np.arccos((normals[..., None] * nw).sum(axis=-2)), 0).max(axis=-1)
        ang = _maxpool3(ang)
        return ang > np.deg2rad(tol)


def postprocess(raw, views, apply_mask=True, mask_edges=True, edge_normal_threshold=5.0,
                edge_depth_threshold=0.03, apply_confidence_mask=False, confidence_percentile=10):
    """inference.py:314-506."""
    mean = torch.tensor(IMG_MEAN).view(1, 1, 1, 3)
    std = torch.tensor(IMG_STD).view(1, 1, 1, 3)
    outs = []
    for r, view in zip(raw, views):
        o = dict(r)
        o["img_no_norm"] = (view["img"].permute(0, 2, 3, 1) * std + mean).clip(0, 1)  # image.py:93-131
        o["depth_z"] = o["pts3d_cam"][..., 2:3]
        o["intrinsics"] = recover_pinhole_intrinsics(o["ray_directions"])
        B = o["cam_trans"].shape[0]
        P = torch.eye(4).unsqueeze(0).repeat(B, 1, 1)
        P[:, :3, :3] = quaternion_to_rotation_matrix(o["cam_quats"])
        P[:, :3, 3] = o["cam_trans"]
        o["camera_poses"] = P
        if apply_mask:
            fm = o["non_ambiguous_mask"].numpy()
            if apply_confidence_mask:
                c = o["conf"]
                thr = torch.quantile(c.reshape(B, -1), confidence_percentile / 100.0, dim=1).view(B, 1, 1)
                fm = fm & (c > thr).numpy()
            if mask_edges:
                pts = o["pts3d"].numpy()
                em = []
                for b in range(B):
                    if fm[b].any():
                        n, nm = points_to_normals(pts[b], fm[b])
                        ne = normals_edge(n, edge_normal_threshold, nm)
                        de = depth_edge(o["depth_z"][b].squeeze(-1).numpy(), edge_depth_threshold, fm[b])
                        em.append(~(de & ne))
                    else:
                        em.append(np.zeros_like(fm[b], dtype=bool))
                fm = fm & np.stack(em, 0)
            fmt = torch.from_numpy(fm).unsqueeze(-1)
            for k in ("pts3d", "pts3d_cam", "depth_along_ray", "depth_z"):
                o[k] = o[k] * fmt
            o["mask"] = fmt
        outs.append(o)
    return outs


def validate_views(views):
    """inference.py:130-219 (exception types and order)."""
    if not views:
        raise ValueError("At least one view must be provided")
    with_pose = []
    for i, v in enumerate(views):
        keys = set(v.keys())
        if "data_norm_type" not in keys:
            raise ValueError(f"View {i} missing required keys: {{'data_norm_type'}}")
        if "intrinsics" in keys and "ray_directions" in keys:
            raise ValueError(f"View {i} contains conflicting keys: ['intrinsics', 'ray_directions'].")
        if i == 0 and "img" not in keys:
            raise ValueError("The First View missing required keys: img")
        if "img" not in keys:
            if "intrinsics" not in keys and "ray_directions" not in keys:
                raise ValueError(f"View {i} without image must provide intrinsics or ray_directions")
            if "camera_poses" not in keys:
                raise ValueError(f"View {i} without image must provide camera_poses")
        if "depth_z" in keys and "intrinsics" not in keys and "ray_directions" not in keys:
            raise ValueError(f"View {i} depth constraint violation")
        if "camera_poses" in keys:
            with_pose.append(i)
    if with_pose and 0 not in with_pose:
        raise ValueError("Camera pose constraint violation")
    return views


def infer(sd, views, apply_mask=True, mask_edges=True, edge_normal_threshold=5.0,
          edge_depth_threshold=0.03, apply_confidence_mask=False, confidence_percentile=10,
          ignore_calibration_inputs=False, ignore_depth_inputs=False, ignore_pose_inputs=False,
          ignore_depth_scale_inputs=False, ignore_pose_scale_inputs=False, taps=None, amp=False,
          aat_opts=None):
    """model.py:2206-2355; fp32 (use_amp=False) unless amp=True (see forward).  Tensors may live on
    any device; post-processing runs on the CPU like the reference's numpy code.  aat_opts: the
    info_sharing module_args that change the arithmetic (see AAT_OPTS)."""
    saved = dict(AAT_OPTS)
    AAT_OPTS.update(aat_opts or {})
    try:
        return _infer(sd, views, apply_mask, mask_edges, edge_normal_threshold, edge_depth_threshold,
                      apply_confidence_mask, confidence_percentile, ignore_calibration_inputs,
                      ignore_depth_inputs, ignore_pose_inputs, ignore_depth_scale_inputs,
                      ignore_pose_scale_inputs, taps, amp)
    finally:
        AAT_OPTS.clear()
        AAT_OPTS.update(saved)


def _infer(sd, views, apply_mask, mask_edges, edge_normal_threshold, edge_depth_threshold,
           apply_confidence_mask, confidence_percentile, ignore_calibration_inputs, ignore_depth_inputs,
           ignore_pose_inputs, ignore_depth_scale_inputs, ignore_pose_scale_inputs, taps, amp):
    with torch.inference_mode():
        pv = preprocess_views(validate_views(views))
        flags = dict(use_calibration=not ignore_calibration_inputs, use_depth=not ignore_depth_inputs,
                     use_pose=not ignore_pose_inputs, use_depth_scale=not ignore_depth_scale_inputs,
                     use_pose_scale=not ignore_pose_scale_inputs)
        raw = forward(sd, pv, flags, taps=taps, amp=amp)
        if raw[0]["pts3d"].device.type != "cpu":
            raw = [{k: v.cpu() for k, v in r.items()} for r in raw]
            pv = [{k: (v.cpu() if torch.is_tensor(v) else v) for k, v in view.items()} for view in pv]
        return postprocess(raw, pv, apply_mask, mask_edges, edge_normal_threshold,
                           edge_depth_threshold, apply_confidence_mask, confidence_percentile)
