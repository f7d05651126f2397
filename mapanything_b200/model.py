"""Host-side mirror of the reference's `MapAnything` class for the inference hot path.

Same constructor / `from_pretrained` / `load_state_dict` / `infer` / `forward` surface, argument
meaning, exception types and prediction-dict keys, shapes and dtypes as
mapanything/models/mapanything/model.py:96-113, 1657, 2206-2355 and
mapanything/utils/inference.py:130-506, but every tensor op runs in hand-written sm_100a kernels
behind the C ABI of libmab200 (include/mab200.h).  PyTorch is used for device memory, streams and
torch.distributed only.  There is no CPU path: without a CUDA device or without libmab200.so this
module raises.
"""
import ctypes as C
import json
import os
from typing import Any, Dict, List, Optional

import torch

from . import _lib, pose_math

# keys a view may carry (inference.py:28-46)
REQUIRED_KEYS = {"data_norm_type"}
CONFLICTING_KEYS = [("intrinsics", "ray_directions")]
_PASS_THROUGH = {"instance", "idx", "true_shape", "data_norm_type"}
_ALL_INPUTS = dict(use_calibration=True, use_depth=True, use_pose=True, use_depth_scale=True,
                   use_pose_scale=True)


def validate_input_views_for_inference(views):
    """Same checks, order and exception type as inference.py:130-219."""
    if not views:
        raise ValueError("At least one view must be provided")
    views_with_poses = []
    for view_idx, view in enumerate(views):
        keys = set(view.keys())
        missing = REQUIRED_KEYS - keys
        if missing:
            raise ValueError(f"View {view_idx} missing required keys: {missing}")
        for conflict in CONFLICTING_KEYS:
            present = [k for k in conflict if k in keys]
            if len(present) > 1:
                raise ValueError(f"View {view_idx} contains conflicting keys: {present}. "
                                 f"Only one of {conflict} can be provided at a time.")
        if view_idx == 0 and "img" not in keys:
            raise ValueError("The First View missing required keys: img")
        if "img" not in keys:
            if "intrinsics" not in keys and "ray_directions" not in keys:
                raise ValueError(f"View {view_idx} without image must provide intrinsics or ray_directions")
            if "camera_poses" not in keys:
                raise ValueError(f"View {view_idx} without image must provide camera_poses")
        if "depth_z" in keys and "intrinsics" not in keys and "ray_directions" not in keys:
            raise ValueError(f"View {view_idx} depth constraint violation: If 'depth_z' is provided, "
                             f"then 'intrinsics' or 'ray_directions' must also be provided.")
        if "camera_poses" in keys:
            views_with_poses.append(view_idx)
    if views_with_poses and 0 not in views_with_poses:
        raise ValueError(f"Camera pose constraint violation: Views {views_with_poses} have camera_poses, "
                         f"but view 0 (reference view) does not.")
    return views


def default_config():
    """Architecture of the production model (configs/model/mapanything.yaml)."""
    return dict(patch=14, enc_dim=1024, enc_depth=24, enc_heads=16, enc_pos_grid=37, aat_dim=768,
                aat_depth=24, aat_heads=12, aat_tap0=11, aat_tap1=17, head_chunk=8,
                aat_ref_view_pe=1, aat_entropy_scaling=0, aat_tap2=0,
                aat_nonref_view_pe=0, aat_all_global=0, head_tf32=1, enc_chunk=64, enc_lanes=1)


def head_ranges(V, stream_chunk, default_chunk):
    """View ranges of the streamed heads.  int: equal ranges of that many views (0: default_chunk).
    Sequence: range sizes in order, the last size repeating (e.g. (4, 2, 1): 4, 2, 1, 1, ...): the tail
    of the call is the device-to-host copy of the LAST range, so the ranges may shrink towards the end."""
    if isinstance(stream_chunk, (tuple, list)):
        sizes = [max(1, int(c)) for c in stream_chunk] or [default_chunk]
    else:
        sizes = [max(1, int(stream_chunk) or default_chunk)]
    out, v0, i = [], 0, 0
    while v0 < V:
        v1 = min(V, v0 + sizes[min(i, len(sizes) - 1)])
        out.append((v0, v1))
        v0, i = v1, i + 1
    return out


def check_state_dict(state_dict, aat_depth=24, ref_view_pe=True, aat_dim=768, view_pe_rows=1):
    """torch.nn.Module.load_state_dict's key / shape bookkeeping for the reference key space
    (synthetic.manifest_for: the reference's own `state_dict()` of the default configuration, aliases
    of shared modules included, re-shaped for the trunk's width and depth).  Returns (missing,
    unexpected, size_mismatch_messages).  A key and its aliases (dense_head.0.* / dpt_feature_head.*,
    SURVEY.md Appendix B) count as one tensor; a checkpoint without the reference-view PE has no
    info_sharing.view_pos_table buffer (aat.py:160-172)."""
    from .synthetic import canonical_key, manifest_for

    expected = {k: tuple(v) for k, v in manifest_for(aat_dim, aat_depth, ref_view_pe, view_pe_rows).items()}
    unexpected = [k for k in state_dict if k not in expected]
    mismatch = [f"size mismatch for {k}: copying a param with shape {tuple(v.shape)} from checkpoint, "
                f"the shape in current model is {expected[k]}."
                for k, v in state_dict.items() if k in expected and tuple(v.shape) != expected[k]]
    have = {canonical_key(k) for k in state_dict if k in expected}
    missing = [k for k in expected if canonical_key(k) == k and k not in have]
    return missing, unexpected, mismatch


class MapAnything(torch.nn.Module):
    """Drop-in for the reference class on the inference path (no trainable state on the torch side:
    the weights live in packed device buffers owned by libmab200)."""

    def __init__(self, name: str = "mapanything", encoder_config: Optional[dict] = None,
                 info_sharing_config: Optional[dict] = None, pred_head_config: Optional[dict] = None,
                 geometric_input_config: Optional[dict] = None, fusion_norm_layer=None,
                 pretrained_checkpoint_path: Optional[str] = None,
                 load_specific_pretrained_submodules: bool = False,
                 specific_pretrained_submodules: Optional[list] = None,
                 torch_hub_force_reload: bool = False, device="cuda", head_chunk: int = 8,
                 head_precision: str = "tf32", enc_chunk: int = 64, enc_lanes: int = 1):
        """Reference constructor arguments (model.py:99-113) plus four of this implementation: `device`,
        `head_chunk` (views per dense-head pass; a workspace knob, results do not depend on it) and
        `enc_chunk` (views per encoder pass; a workspace knob like head_chunk: the encoder's activations are
        held for enc_chunk views at a time, which is what lets thousands of views fit one GPU) and
        `head_precision`: "tf32" (default) runs the DPT feature head and the regressor with fp32
        activations and TF32 contractions, the reference's own arithmetic for that region (autocast off,
        model.py:1774); "bf16" is the faster opt-in with bf16 operands."""
        super().__init__()
        if head_precision not in ("tf32", "bf16"):
            raise ValueError(f"head_precision must be 'tf32' or 'bf16', got {head_precision!r}")
        self.name = name
        self.encoder_config = encoder_config or {"data_norm_type": "dinov2", "name": "dinov2_large"}
        self.info_sharing_config = info_sharing_config or {}
        self.pred_head_config = pred_head_config or {}
        self.geometric_input_config = geometric_input_config or {}
        self.data_norm_type = self.encoder_config.get("data_norm_type", "dinov2")
        self._check_supported()
        if not torch.cuda.is_available():
            raise RuntimeError("mapanything_b200 needs a CUDA device (sm_100a); there is no CPU path")
        self._device = torch.device(device if device != "cuda" else f"cuda:{torch.cuda.current_device()}")
        self._lib = _lib.load()
        cfg = default_config()
        margs = self.info_sharing_config.get("module_args", {})
        if "depth" in margs:
            cfg["aat_depth"] = int(margs["depth"])
        if "indices" in margs:  # two or three intermediate features (model.py:324-327)
            taps = [int(i) for i in margs["indices"]]
            cfg["aat_tap0"], cfg["aat_tap1"], cfg["aat_tap2"] = (taps + [0])[:3]
        # aat_ifr_48_layers*.yaml: width 1024, 16 heads
        cfg["aat_dim"] = int(margs.get("dim", 768))
        cfg["aat_heads"] = int(margs.get("num_heads", cfg["aat_dim"] // 64))
        # info_sharing variants: aat_ifr_24_layers_no_ref_view.yaml / aat_ifr_24_layers_escaling.yaml
        cfg["aat_ref_view_pe"] = int(bool(margs.get("distinguish_ref_and_non_ref_views", True)))
        cfg["aat_entropy_scaling"] = int(bool(margs.get("use_entropy_scaling", False)))
        # use_pe_for_non_reference_views: "Only used if distinguish_ref_and_non_ref_views flag is True" (aat.py:64)
        cfg["aat_nonref_view_pe"] = int(bool(margs.get("use_pe_for_non_reference_views", False)) and
                                        bool(cfg["aat_ref_view_pe"]))
        self._max_views_for_pe = int(margs.get("max_num_views_for_pe", 1000))
        self._rand_view_pe = bool(margs.get("use_rand_idx_pe_for_non_reference_views", True))
        self._check_view_count = bool(margs.get("use_pe_for_non_reference_views", False))
        if self.info_sharing_config.get("model_type", "alternating_attention") == "global_attention":
            # gat_ifr_24_layers*.yaml: every block is global and every view is always encoded with a row of
            # the [max_num_views, dim] table (global_attention_transformer.py:543-560)
            cfg["aat_all_global"] = cfg["aat_ref_view_pe"] = cfg["aat_nonref_view_pe"] = 1
            self._max_views_for_pe = int(margs["max_num_views"])
            self._rand_view_pe = bool(margs["use_rand_idx_pe_for_non_reference_views"])
            self._check_view_count = True
        self._pinned_view_pe_rows = None
        cfg["head_chunk"] = int(head_chunk)
        cfg["head_tf32"] = int(head_precision == "tf32")
        cfg["enc_chunk"] = int(enc_chunk)
        cfg["enc_lanes"] = int(os.environ.get("MAB_ENC_LANES", enc_lanes))  # env: A/B runs (profiles/)
        self.head_precision = head_precision
        self._cfg = _lib.MabConfig(**cfg)
        self._handle = C.c_void_p()
        _lib.check(self._lib.mab_create(C.byref(self._cfg), self._device.index or 0, C.byref(self._handle)),
                   "mab_create")
        self._finalized = False
        self._workspace = None
        self._post_ws = None
        self._host_pool = None
        self._copy_stream = None
        self._ring_bufs = None
        self._kv_all = None
        self._comm = None   # (ncclComm_t, rank, world) given to the library for the one-call sharded forward
        if pretrained_checkpoint_path is not None:
            ckpt = torch.load(pretrained_checkpoint_path, map_location="cpu", weights_only=False)
            self.load_state_dict(ckpt["model"] if "model" in ckpt else ckpt)

    # ------------------------------------------------------------------ construction helpers
    def _check_supported(self):
        ic = self.info_sharing_config
        if ic and ic.get("model_type", "alternating_attention") not in ("alternating_attention", "global_attention"):
            raise NotImplementedError("the alternating- and global-attention info-sharing models are built; "
                                      "cross_attention (cat_ifr_dust3r.yaml) is a two-view decoder of its own")
        if ic and ic.get("model_type") == "global_attention":
            for need in ("max_num_views", "use_rand_idx_pe_for_non_reference_views"):
                if need not in ic.get("module_args", {}):  # positional in global_attention_transformer.py:350-356
                    raise TypeError(f"MultiViewGlobalAttentionTransformerIFR.__init__() missing 1 required "
                                    f"positional argument: '{need}'")
        cpe = (ic or {}).get("custom_positional_encoding")
        if isinstance(cpe, str):  # model.py:259-267: the reference implements none of the named encodings
            raise ValueError(f"Invalid custom_positional_encoding: {cpe}. None implemented.")
        if cpe is not None:
            # a callable reaches aat.py:612-616, where the position tensor is concatenated with the [None]
            # of the scale token (TypeError in the reference): no MapAnything configuration runs with one
            raise NotImplementedError("callable positional encodings do not run in the reference with the "
                                      "scale token (aat.py:612-616); mab_rope2d is the RoPE2D kernel on its own")
        margs = (ic or {}).get("module_args", {})
        dim = int(margs.get("dim", 768))
        if dim not in (768, 1024) or int(margs.get("num_heads", dim // 64)) * 64 != dim:
            raise NotImplementedError("the info-sharing transformer is built for widths 768 and 1024 with "
                                      "64-wide heads (aat_ifr_24_layers*.yaml, aat_ifr_48_layers*.yaml)")
        if len(margs.get("indices", (11, 17))) not in (2, 3):
            raise ValueError("Invalid number of indices provided for info sharing feature returner. "
                             "Please provide 2 or 3 indices.")  # model.py:328-331
        pc = self.pred_head_config
        if pc and pc.get("type", "dpt+pose") != "dpt+pose":
            raise NotImplementedError("only the dpt+pose head is built")
        # the fused regressor epilogue is the production adaptor (pred_head/adaptor_config/
        # raydirs_depth_pose_confidence_mask_scale.yaml): anything else must be refused, not mis-computed
        inf = float("inf")
        want = {
            None: dict(adaptor_type="raydirs+depth+pose+confidence+mask"),
            "feature_head": dict(feature_dim=256, hooks=[0, 1, 2, 3]),
            "regressor_head": dict(output_dim=6),
            "pose_head": dict(num_resconv_block=2, rot_representation_dim=4),
            "scale_head": dict(output_dim=1),
            "dpt_adaptor": dict(ray_directions_mode="linear", ray_directions_normalize_to_unit_sphere=True,
                                ray_directions_normalize_to_unit_image_plane=False, ray_directions_vmin=-inf,
                                ray_directions_vmax=inf, ray_directions_clamp_min_of_z_dir=False, depth_mode="exp",
                                depth_vmin=0, depth_vmax=inf, confidence_type="exp", confidence_vmin=1,
                                confidence_vmax=inf),
            "pose_adaptor": dict(cam_trans_mode="linear", cam_trans_vmin=-inf, cam_trans_vmax=inf,
                                 quaternions_mode="linear", quaternions_normalize=True, quaternions_vmin=-inf,
                                 quaternions_vmax=inf),
            "scale_adaptor": dict(mode="exp", vmin=1e-8, vmax=inf),
        }
        for section, fields in want.items():
            got = pc if section is None else pc.get(section) or {}
            for k, v in fields.items():
                if k in got and (list(got[k]) if isinstance(v, list) else got[k]) != v:
                    where = f"pred_head_config[{section!r}]" if section else "pred_head_config"
                    raise NotImplementedError(f"{where}[{k!r}] = {got[k]!r}: the built head is the production "
                                              f"configuration ({k} = {v!r})")
        ec = self.encoder_config
        if ec.get("name", "dinov2_large") != "dinov2_large" or ec.get("with_registers", False):
            raise NotImplementedError("only the dinov2_large encoder (ViT-L/14, no registers) is built")
        if self.data_norm_type != "dinov2":
            raise NotImplementedError("the encoder takes dinov2-normalised images")

    @classmethod
    def from_pretrained(cls, path: str, device="cuda", **kwargs):
        """Local-directory flavour of PyTorchModelHubMixin.from_pretrained (model.py:96):
        `config.json` -> constructor kwargs, `model.safetensors` (or `model.pth`) -> state dict."""
        if not os.path.isdir(path):
            raise FileNotFoundError(f"{path}: hub downloads need network access; pass a local directory "
                                    f"holding config.json and model.safetensors")
        with open(os.path.join(path, "config.json")) as f:
            config = json.load(f)
        config = {k: v for k, v in config.items() if k in (
            "name", "encoder_config", "info_sharing_config", "pred_head_config", "geometric_input_config")}
        model = cls(device=device, **config, **kwargs)
        st_path = os.path.join(path, "model.safetensors")
        if os.path.exists(st_path):
            from safetensors.torch import load_file
            sd = load_file(st_path)
        else:
            sd = torch.load(os.path.join(path, "model.pth"), map_location="cpu", weights_only=False)
            sd = sd.get("model", sd)
        model.load_state_dict(sd)
        return model

    @property
    def device(self):
        return self._device

    def to(self, device=None, *a, **k):
        if device is not None and torch.device(device).type != "cuda":
            raise RuntimeError("mapanything_b200 runs on CUDA devices only")
        return self

    def load_state_dict(self, state_dict, strict: bool = True):
        """Accepts the reference key space unchanged, aliases included."""
        if self._finalized:
            raise RuntimeError("weights are already finalized for this instance")
        missing, unexpected, mismatch = check_state_dict(
            state_dict, self._cfg.aat_depth, bool(self._cfg.aat_ref_view_pe), self._cfg.aat_dim,
            self._max_views_for_pe if self._cfg.aat_nonref_view_pe else 1)
        errors = list(mismatch)  # a size mismatch is an error with strict=False as well (torch)
        if strict and unexpected:
            errors.insert(0, "Unexpected key(s) in state_dict: " + ", ".join(f'"{k}"' for k in unexpected) + ". ")
        if missing:  # every weight is needed to run: no partially initialised model on this path
            errors.insert(0, "Missing key(s) in state_dict: " + ", ".join(f'"{k}"' for k in missing) + ". ")
        if errors:
            raise RuntimeError("Error(s) in loading state_dict for MapAnything:\n\t" + "\n\t".join(errors))
        with torch.cuda.device(self._device):
            for key, t in state_dict.items():
                if key in unexpected:
                    continue
                t = t.detach().to(dtype=torch.float32).contiguous()
                shape = (C.c_int64 * max(t.dim(), 1))(*(list(t.shape) or [1]))
                _lib.check(self._lib.mab_set_weight(self._handle, key.encode(), t.data_ptr(), 0,
                                                    max(t.dim(), 1), shape), f"mab_set_weight({key})")
            _lib.check(self._lib.mab_finalize_weights(self._handle, self._stream_ptr()),
                       "mab_finalize_weights")
        self._finalized = True
        return torch.nn.modules.module._IncompatibleKeys([], unexpected)

    def __del__(self):
        try:
            if getattr(self, "_handle", None):
                self._lib.mab_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    # ------------------------------------------------------------------ plumbing
    def _stream_ptr(self):
        return C.c_void_p(torch.cuda.current_stream(self._device).cuda_stream)

    def _get_workspace(self, nbytes):
        if self._workspace is None or self._workspace.numel() < nbytes:
            self._workspace = None
            self._workspace = torch.empty(int(nbytes), dtype=torch.uint8, device=self._device)
        return self._workspace

    def _pack_images(self, views):
        shapes = {tuple(v["img"].shape) for v in views}
        assert len(shapes) == 1, "All views must have the same image shape"  # model.py:1666-1667
        B, Cc, H, W = views[0]["img"].shape
        assert Cc == 3, "Input must have 3 channels"
        assert H % 14 == 0 and W % 14 == 0, "Input shape must be divisible by patch size: 14"
        assert B == 1, "a batch of B scenes is split into scenes by infer() / forward()"
        imgs = torch.empty((len(views), 3, H, W), dtype=torch.float32, device=self._device)
        for i, v in enumerate(views):
            imgs[i].copy_(v["img"][0], non_blocking=True)
        return imgs, H, W

    def _pack_geometry(self, views, flags, H, W, view_lo=0, total_views=None, default_metric=True):
        """a1 + the host part of a6: pre-process the optional geometric inputs of this rank's views
        (inference.py:222-311) into the stacked device buffers `mab_inputs` takes, with the per-view
        availability masks of model.py:1336-1369 made deterministic by `flags`
        (model.py:2154-2197).  Returns None when nothing geometric is used (image-only path).

        Per-pixel work (rays from intrinsics, ray normalisation, z-depth -> depth along ray) runs
        in libmab200 kernels; the pose algebra (views x 7 floats) is host code (pose_math.py)."""
        V, dev, lib, st = len(views), self._device, self._lib, self._stream_ptr()
        if not (flags["use_calibration"] or flags["use_depth"] or flags["use_pose"]):
            return None
        rays = torch.zeros((V, H, W, 3), dtype=torch.float32, device=dev)
        depth = torch.zeros((V, H, W), dtype=torch.float32, device=dev)
        # is_metric_scale: infer()'s pre-processing defaults it to True (inference.py:298-306); the raw
        # forward() uses zeros when the key is absent (model.py:1044-1053)
        has_ray, has_depth, has_cam, metric = [0] * V, [0] * V, [0] * V, [int(default_metric)] * V
        quats, trans = [None] * V, [None] * V
        npix = H * W
        with torch.cuda.device(dev):
            for i, v in enumerate(views):
                if "intrinsics" in v:
                    K = v["intrinsics"].to(dev, torch.float32).reshape(-1, 3, 3)[:1].contiguous()
                    _lib.check(lib.mab_rays_from_intrinsics(K.data_ptr(), 1, H, W, rays[i].data_ptr(), st))
                    has_ray[i] = 1
                elif "ray_directions" in v or "ray_directions_cam" in v:
                    key = "ray_directions" if "ray_directions" in v else "ray_directions_cam"
                    r = v[key].to(dev, torch.float32).reshape(-1, H, W, 3)[:1].contiguous()
                    if key == "ray_directions":
                        _lib.check(lib.mab_normalize_rays(r.data_ptr(), npix, rays[i].data_ptr(), st))
                    else:  # already pre-processed (forward() on the output of preprocess_input_views)
                        rays[i].copy_(r[0])
                    has_ray[i] = 1
                if "depth_z" in v:
                    z = v["depth_z"].to(dev, torch.float32).reshape(-1, H, W)[:1].contiguous()
                    _lib.check(lib.mab_depth_z_to_along_ray(z.data_ptr(), rays[i].data_ptr(), npix,
                                                            depth[i].data_ptr(), st))
                    has_depth[i] = 1
                elif "depth_along_ray" in v:
                    depth[i].copy_(v["depth_along_ray"].to(dev, torch.float32).reshape(-1, H, W)[0])
                    has_depth[i] = 1
                if "camera_poses" in v:
                    q, t = pose_math.split_camera_poses(v["camera_poses"], view_lo + i)
                    quats[i], trans[i], has_cam[i] = q, t, 1
                elif "camera_pose_quats" in v and "camera_pose_trans" in v:
                    quats[i], trans[i], has_cam[i] = v["camera_pose_quats"], v["camera_pose_trans"], 1
                if "is_metric_scale" in v:  # default True, inference.py:298-306
                    metric[i] = int(bool(torch.as_tensor(v["is_metric_scale"]).reshape(-1)[0]))
        ray_mask = [int(flags["use_calibration"] and has_ray[i]) for i in range(V)]
        depth_mask = [int(flags["use_depth"] and has_depth[i]) for i in range(V)]
        cam_mask = [int(flags["use_pose"] and has_cam[i]) for i in range(V)]

        # poses relative to the scene's view 0 and the scene-wide translation norm (model.py:792-896,
        # geometry.py:1629-1666).  A view-sharded scene exchanges the handful of pose floats first.
        local = [(cam_mask[i], None if not cam_mask[i] else
                  (torch.as_tensor(quats[i]).detach().float().cpu().reshape(-1, 4)[:1],
                   torch.as_tensor(trans[i]).detach().float().cpu().reshape(-1, 3)[:1])) for i in range(V)]
        total = int(total_views or V)
        scene = local
        if total > V:
            import torch.distributed as dist
            gathered = [None] * dist.get_world_size()
            dist.all_gather_object(gathered, local)
            scene = [e for part in gathered for e in part]
        q_rel = torch.tensor([0.0, 0.0, 0.0, 1.0]).repeat(len(scene), 1)
        t_rel = torch.zeros(len(scene), 3)
        pose_norm = 1.0
        posed = [i for i, (m, _) in enumerate(scene) if m]
        if posed:
            if not scene[0][0]:
                raise ValueError("camera poses are given but the reference view (view 0) has none")
            q_all = torch.cat([scene[i][1][0] for i in posed], 0)
            t_all = torch.cat([scene[i][1][1] for i in posed], 0)
            q, t = pose_math.relative_to_first(scene[0][1][0], scene[0][1][1], q_all, t_all)
            q_rel[posed], t_rel[posed] = q, t
            pose_norm = pose_math.translation_norm(t_rel)
        sl = slice(view_lo, view_lo + V) if total > V else slice(0, V)

        def u8(x):
            return torch.tensor(x, dtype=torch.uint8).to(dev, non_blocking=True)

        geo = dict(
            ray_dirs=rays if any(ray_mask) else None,
            depth_along_ray=depth if any(depth_mask) else None,
            cam_quats=q_rel[sl].contiguous().to(dev) if any(cam_mask) else None,
            cam_trans=t_rel[sl].contiguous().to(dev) if any(cam_mask) else None,
            ray_mask=u8(ray_mask), depth_mask=u8(depth_mask), cam_mask=u8(cam_mask),
            depth_metric_mask=u8([int(metric[i] and flags["use_depth_scale"]) for i in range(V)]),
            pose_metric_mask=u8([int(metric[i] and flags["use_pose_scale"]) for i in range(V)]),
            pose_trans_norm=float(pose_norm),
        )
        if not (any(ray_mask) or any(depth_mask) or any(cam_mask)):
            return None
        return geo

    def _input_struct(self, imgs, H, W, geo=None):
        inp = _lib.MabInputs(views=imgs.shape[0], H=H, W=W, images=imgs.data_ptr())
        if geo:
            for k in ("ray_dirs", "depth_along_ray", "cam_quats", "cam_trans", "ray_mask", "depth_mask",
                      "cam_mask", "depth_metric_mask", "pose_metric_mask"):
                if geo[k] is not None:
                    setattr(inp, k, geo[k].data_ptr())
            inp.pose_trans_norm = geo["pose_trans_norm"]
        return inp

    def _alloc_outputs(self, V, H, W, want_logits=False):
        dev = self._device
        o = dict(
            pts3d=torch.empty((V, H, W, 3), dtype=torch.float32, device=dev),
            pts3d_cam=torch.empty((V, H, W, 3), dtype=torch.float32, device=dev),
            ray_directions=torch.empty((V, H, W, 3), dtype=torch.float32, device=dev),
            depth_along_ray=torch.empty((V, H, W, 1), dtype=torch.float32, device=dev),
            conf=torch.empty((V, H, W), dtype=torch.float32, device=dev),
            mask_logits=torch.empty((V, H, W), dtype=torch.float32, device=dev),
            non_ambiguous_mask=torch.empty((V, H, W), dtype=torch.bool, device=dev),
            pose=torch.empty((V, 8), dtype=torch.float32, device=dev),
            metric_scale=torch.empty((1, 1), dtype=torch.float32, device=dev),
        )
        if want_logits:
            o["logits6"] = torch.empty((V, H, W, 6), dtype=torch.float32, device=dev)
        return o

    @staticmethod
    def _out_struct(o):
        return _lib.MabOutputs(**{k: (o[k].data_ptr() if k in o else None) for k, _ in _lib.MabOutputs._fields_})

    def _run_forward(self, imgs, H, W, want_logits=False, view_lo=0, total_views=None, geo=None,
                     on_views=None, stream_chunk=0):
        """One whole forward of this rank's views: returns the raw output tensors (views stacked).

        Single GPU (total_views == views): one `mab_forward` call.  View-sharded
        (total_views > views, torch.distributed initialised, rank r holds views [view_lo, ...)): one
        `mab_forward_sharded` call -- the library exchanges the K|V rows of the global layers itself over
        an NCCL communicator of its own (all-to-all gather overlapped with the local block) and broadcasts
        the metric scale from the rank that owns the scale token; MAB_EXCHANGE=ring / MAB_EXCHANGE_HOST=1
        select the per-layer schedules sequenced here through torch.distributed instead
        (parallel.ring_pass / parallel.gather_pass over the staged C ABI)."""
        V = imgs.shape[0]
        total = int(total_views or V)
        with torch.cuda.device(self._device):
            nbytes = self._lib.mab_workspace_bytes(self._handle, V, total, H, W)
            if nbytes == 0:
                raise RuntimeError("mab_workspace_bytes rejected the problem shape")
            ws = self._get_workspace(nbytes)
            self._set_view_pe_rows(view_lo, V, total)
            outs = self._alloc_outputs(V, H, W, want_logits)
            inp = self._input_struct(imgs, H, W, geo)
            ost = self._out_struct(outs)
            if on_views is not None:
                # streamed heads: trunk for all views, then the heads `stream_chunk` views at a time,
                # handing every finished range to the caller (post-processing + copy-out overlap the
                # heads of the later views)
                if total == V:
                    self._trunk_single(inp, outs, ws)
                else:
                    self._forward_sharded(inp, ost, outs, ws, V, H, W, view_lo, total, heads=False)
                for v0, v1 in self._head_ranges(V, stream_chunk):
                    _lib.check(self._lib.mab_stage_heads_range(self._handle, C.byref(inp), C.byref(ost), v0, v1,
                                                               self._stream_ptr()), "mab_stage_heads_range")
                    on_views(outs, v0, v1)
            elif total == V:
                _lib.check(self._lib.mab_forward(self._handle, C.byref(inp), C.byref(ost), ws.data_ptr(),
                                                 ws.numel(), self._stream_ptr()), "mab_forward")
            else:
                self._forward_sharded(inp, ost, outs, ws, V, H, W, view_lo, total)
        return outs

    def _draw_view_pe_rows(self, total, sharded=False):
        """view_pos_table row of every view of a scene (aat.py:618-645): 0 for the reference view, then
        torch.randint(1, max_num_views_for_pe) from the global CPU generator like the reference (default)
        or 1..V-1.  A view-SHARDED call (this rank holds fewer views than the scene has) takes rank 0's
        draw; ranks that run independent scenes under an initialised process group draw locally -- no
        hidden collective outside the sharded path."""
        if self._rand_view_pe:
            rows = torch.randint(low=1, high=self._max_views_for_pe, size=(total - 1,))
        else:
            rows = torch.arange(1, total)
        rows = torch.cat([torch.zeros(1, dtype=torch.int64), rows.to(torch.int64)])
        import torch.distributed as dist
        if (sharded and self._rand_view_pe and dist.is_available() and dist.is_initialized()
                and dist.get_world_size() > 1):
            dev_rows = rows.to(self._device)
            dist.broadcast(dev_rows, src=0)
            rows = dev_rows.cpu()
        return rows

    def _set_view_pe_rows(self, view_lo, V, total):
        if self._check_view_count:  # aat.py:554-557
            assert total <= self._max_views_for_pe, f"Expected less than {self._max_views_for_pe} views, got {total}"
        if not self._cfg.aat_nonref_view_pe:
            return
        rows = (self._pinned_view_pe_rows if self._pinned_view_pe_rows is not None
                else self._draw_view_pe_rows(total, sharded=total > V))
        local = (C.c_int * V)(*[int(r) for r in rows[view_lo:view_lo + V]])
        _lib.check(self._lib.mab_set_view_pe_rows(self._handle, local, V), "mab_set_view_pe_rows")

    def _head_ranges(self, V, stream_chunk):
        return head_ranges(V, stream_chunk, self._cfg.head_chunk)

    def _trunk_single(self, inp, outs, ws):
        """Encoder + info sharing + scale head of a whole scene on this GPU through the staged ABI
        (what mab_forward does internally), leaving the per-view heads to the caller."""
        lib, hd, st = self._lib, self._handle, self._stream_ptr()
        _lib.check(lib.mab_stage_encode(hd, C.byref(inp), inp.views, ws.data_ptr(), ws.numel(), 3, st),
                   "mab_stage_encode")
        both = _lib.RING_FIRST | _lib.RING_LAST
        for layer in range(self._cfg.aat_depth):
            _lib.check(lib.mab_stage_aat_qkv(hd, layer, None, None, st), "mab_stage_aat_qkv")
            _lib.check(lib.mab_stage_aat_attend(hd, layer, None, 0, both, st), "mab_stage_aat_attend")
            _lib.check(lib.mab_stage_aat_finish(hd, layer, st), "mab_stage_aat_finish")
        _lib.check(lib.mab_stage_scale_head(hd, outs["metric_scale"].data_ptr(), st), "mab_stage_scale_head")

    def _forward_sharded(self, inp, ost, outs, ws, V, H, W, view_lo, total, heads=True):
        import torch.distributed as dist

        from . import parallel

        lib, hd = self._lib, self._handle
        world, rank = dist.get_world_size(), dist.get_rank()
        assert parallel.shard_views(total, world, rank) == (view_lo, view_lo + V), "views do not match the shard table"
        T = (H // 14) * (W // 14)
        width = 3 * self._cfg.aat_dim
        rows = parallel.rows_per_rank(total, world, T)
        # exchange step of the global layers: all-to-all gather of K|V (default) or the ring; the
        # gather needs the whole scene's K|V resident on every rank (1.6 GB at 512 views)
        kv_bytes = sum(rows) * 2 * self._cfg.aat_dim * 2
        mode = os.environ.get("MAB_EXCHANGE", "gather")
        if mode == "gather" and kv_bytes > int(os.environ.get("MAB_GATHER_MAX_BYTES", 24 << 30)):
            mode = "ring"
        if mode == "gather" and (self._kv_all is None or self._kv_all.shape[0] != sum(rows)):
            self._kv_all = torch.empty((sum(rows), 2 * self._cfg.aat_dim), dtype=torch.bfloat16,
                                       device=self._device)
        if mode == "gather" and os.environ.get("MAB_EXCHANGE_HOST", "0") != "1":
            # default: the whole sharded forward is ONE library call (mab_forward_sharded): the K|V exchange
            # of the global layers is issued by the library on its own stream through an NCCL communicator
            # created once for the ranks of the default group (MAB_EXCHANGE_HOST=1: the per-layer exchange
            # sequenced from here through torch.distributed, the round-1 path)
            if self._comm is None:
                from . import nccl_comm
                comm, r, w, path = nccl_comm.create(self._device)
                _lib.check(lib.mab_set_comm(hd, comm, r, w, path.encode()), "mab_set_comm")
                self._comm = (comm, r, w)
            _lib.check(lib.mab_forward_sharded(hd, C.byref(inp), total, C.byref(ost), ws.data_ptr(), ws.numel(),
                                               self._kv_all.data_ptr(), self._kv_all.numel() * 2, int(heads),
                                               self._stream_ptr()), "mab_forward_sharded")
            return
        if mode == "ring" and (self._ring_bufs is None or self._ring_bufs[0].shape[0] < max(rows)):
            self._ring_bufs = [torch.empty((max(rows), width), dtype=torch.bfloat16, device=self._device)
                               for _ in range(2)]
        own_first = 3 if view_lo == 0 else 0
        _lib.check(lib.mab_stage_encode(hd, C.byref(inp), total, ws.data_ptr(), ws.numel(), own_first,
                                        self._stream_ptr()), "mab_stage_encode")
        qkv_ptr, n_rows = C.c_void_p(), C.c_int()
        both = _lib.RING_FIRST | _lib.RING_LAST
        for layer in range(self._cfg.aat_depth):
            _lib.check(lib.mab_stage_aat_qkv(hd, layer, C.byref(qkv_ptr), C.byref(n_rows), self._stream_ptr()),
                       "mab_stage_aat_qkv")
            if self._cfg.aat_all_global or layer % 2 == 0:
                assert n_rows.value == rows[rank]
                off = qkv_ptr.value - ws.data_ptr()
                local = ws[off: off + n_rows.value * width * 2].view(torch.bfloat16).view(n_rows.value, width)

                def attend(block, n, step, src, layer=layer):
                    flags = (_lib.RING_FIRST if step == 0 else 0) | (_lib.RING_LAST if step == world - 1 else 0)
                    _lib.check(lib.mab_stage_aat_attend(hd, layer, block.data_ptr(), n, flags, self._stream_ptr()),
                               "mab_stage_aat_attend")

                def flags_of(first, last):
                    return (_lib.RING_FIRST if first else 0) | (_lib.RING_LAST if last else 0)

                def export_local(dst):
                    _lib.check(lib.mab_stage_aat_export_kv(hd, dst.data_ptr(), self._stream_ptr()),
                               "mab_stage_aat_export_kv")

                def attend_local(first, last, layer=layer):
                    _lib.check(lib.mab_stage_aat_attend(hd, layer, None, 0, flags_of(first, last),
                                                        self._stream_ptr()), "mab_stage_aat_attend")

                def attend_kv(block, n, first, last, layer=layer):
                    _lib.check(lib.mab_stage_aat_attend_kv(hd, layer, block.data_ptr(), n, flags_of(first, last),
                                                           self._stream_ptr()), "mab_stage_aat_attend_kv")

                if mode == "gather":
                    parallel.gather_pass(rows, self._kv_all, export_local, attend_local, attend_kv)
                else:
                    parallel.ring_pass(local, rows, self._ring_bufs, attend)
            else:
                _lib.check(lib.mab_stage_aat_attend(hd, layer, None, 0, both, self._stream_ptr()),
                           "mab_stage_aat_attend")
            _lib.check(lib.mab_stage_aat_finish(hd, layer, self._stream_ptr()), "mab_stage_aat_finish")
        if view_lo == 0:
            _lib.check(lib.mab_stage_scale_head(hd, outs["metric_scale"].data_ptr(), self._stream_ptr()),
                       "mab_stage_scale_head")
        dist.broadcast(outs["metric_scale"], src=0)
        if heads:
            _lib.check(lib.mab_stage_heads(hd, C.byref(inp), C.byref(ost), outs["metric_scale"].data_ptr(),
                                           self._stream_ptr()), "mab_stage_heads")

    def _alloc_post(self, V, H, W):
        dev = self._device
        return dict(
            img_no_norm=torch.empty((V, H, W, 3), dtype=torch.float32, device=dev),
            depth_z=torch.empty((V, H, W, 1), dtype=torch.float32, device=dev),
            intrinsics=torch.empty((V, 3, 3), dtype=torch.float32, device=dev),
            camera_poses=torch.empty((V, 4, 4), dtype=torch.float32, device=dev),
            cam_trans=torch.empty((V, 3), dtype=torch.float32, device=dev),
            cam_quats=torch.empty((V, 4), dtype=torch.float32, device=dev),
            mask=torch.empty((V, H, W, 1), dtype=torch.bool, device=dev),
        )

    def _run_postprocess(self, imgs, outs, H, W, apply_mask=True, mask_edges=True, edge_normal_threshold=5.0,
                         edge_depth_threshold=0.03, apply_confidence_mask=False, confidence_percentile=10,
                         post=None, view_lo=0, view_hi=None):
        """inference.py:314-506 on the device: all views of this rank at once, or the range
        [view_lo, view_hi) of them (every step is per-view independent)."""
        dev = self._device
        view_hi = imgs.shape[0] if view_hi is None else view_hi
        if post is None:
            post = self._alloc_post(imgs.shape[0], H, W)
        V = view_hi - view_lo
        sl = slice(view_lo, view_hi)
        with torch.cuda.device(dev):
            need = V * H * W * 18 + V * 8 + 1024
            if self._post_ws is None or self._post_ws.numel() < need:
                self._post_ws = torch.empty(need, dtype=torch.uint8, device=dev)
            inp = _lib.MabInputs(views=V, H=H, W=W, images=imgs[sl].data_ptr())
            ost = _lib.MabOutputs(**{k: ((outs[k] if k == "metric_scale" else outs[k][sl]).data_ptr() if k in outs else None)
                                     for k, _ in _lib.MabOutputs._fields_})
            pst = _lib.MabPostOutputs(**{k: post[k][sl].data_ptr() for k, _ in _lib.MabPostOutputs._fields_})
            opt = _lib.MabPostOptions(int(apply_mask), int(mask_edges), float(edge_normal_threshold),
                                      float(edge_depth_threshold), int(apply_confidence_mask),
                                      float(confidence_percentile))
            _lib.check(self._lib.mab_postprocess(self._handle, C.byref(inp), C.byref(ost), C.byref(pst),
                                                 C.byref(opt), self._post_ws.data_ptr(), self._post_ws.numel(),
                                                 self._stream_ptr()), "mab_postprocess")
        return post

    @torch.inference_mode()
    def run_resident(self, imgs, H, W, view_range=None, total_views=None, **post_kw):
        """Forward + post-processing on an image stack that is already resident on the device
        ((views,3,H,W) fp32).  Returns the stacked raw and post-processed tensors (no dicts)."""
        lo = view_range[0] if view_range else 0
        outs = self._run_forward(imgs, H, W, view_lo=lo, total_views=total_views)
        return outs, self._run_postprocess(imgs, outs, H, W, **post_kw)

    def _raw_dicts(self, outs, V):
        scale = outs["metric_scale"]
        cam_trans = outs["pose"][:, 0:3] * scale  # (V,3)
        cam_quats = outs["pose"][:, 3:7]
        res = []
        for i in range(V):
            res.append({
                "pts3d": outs["pts3d"][i:i + 1],
                "pts3d_cam": outs["pts3d_cam"][i:i + 1],
                "ray_directions": outs["ray_directions"][i:i + 1],
                "depth_along_ray": outs["depth_along_ray"][i:i + 1],
                "cam_trans": cam_trans[i:i + 1],
                "cam_quats": cam_quats[i:i + 1].contiguous(),
                "metric_scaling_factor": scale,
                "conf": outs["conf"][i:i + 1],
                "non_ambiguous_mask": outs["non_ambiguous_mask"][i:i + 1],
                "non_ambiguous_mask_logits": outs["mask_logits"][i:i + 1],
            })
        return res

    # prediction key -> (source dict, key there)
    _PRED_KEYS = (("pts3d", "raw", "pts3d"), ("pts3d_cam", "raw", "pts3d_cam"),
                  ("ray_directions", "raw", "ray_directions"), ("depth_along_ray", "raw", "depth_along_ray"),
                  ("cam_trans", "post", "cam_trans"), ("cam_quats", "post", "cam_quats"),
                  ("conf", "raw", "conf"), ("non_ambiguous_mask", "raw", "non_ambiguous_mask"),
                  ("non_ambiguous_mask_logits", "raw", "mask_logits"), ("img_no_norm", "post", "img_no_norm"),
                  ("depth_z", "post", "depth_z"), ("intrinsics", "post", "intrinsics"),
                  ("camera_poses", "post", "camera_poses"), ("mask", "post", "mask"))

    # per-view shape of every prediction (-1 = H, -2 = W) and its dtype
    _PRED_SHAPES = {"pts3d": ((-1, -2, 3), torch.float32), "pts3d_cam": ((-1, -2, 3), torch.float32),
                    "ray_directions": ((-1, -2, 3), torch.float32), "depth_along_ray": ((-1, -2, 1), torch.float32),
                    "cam_trans": ((3,), torch.float32), "cam_quats": ((4,), torch.float32),
                    "conf": ((-1, -2), torch.float32), "non_ambiguous_mask": ((-1, -2), torch.bool),
                    "non_ambiguous_mask_logits": ((-1, -2), torch.float32), "img_no_norm": ((-1, -2, 3), torch.float32),
                    "depth_z": ((-1, -2, 1), torch.float32), "intrinsics": ((3, 3), torch.float32),
                    "camera_poses": ((4, 4), torch.float32), "mask": ((-1, -2, 1), torch.bool)}

    def _infer_to_host(self, imgs, H, W, view_lo, total_views, geo, post_args, stream_chunk,
                       reuse_host_buffers=False):
        """infer(host_outputs=True): heads in ranges of `stream_chunk` views; post-processing of a
        finished range on the compute stream, its device-to-host copies on a second stream."""
        V, dev = imgs.shape[0], self._device
        apply_mask = post_args[0]
        post = self._alloc_post(V, H, W)
        key = (V, H, W)
        if not reuse_host_buffers or self._host_pool is None or self._host_pool[0] != key:
            # fresh, caller-owned pinned tensors (the reference returns fresh tensors); torch's
            # pinned-memory caching allocator recycles the blocks of results the caller has dropped
            bufs = {}
            for name, src, k in self._PRED_KEYS:
                shape, dtype = self._PRED_SHAPES[name]
                bufs[name] = torch.empty((V,) + tuple(x if x >= 0 else (H, W)[-x - 1] for x in shape), dtype=dtype,
                                         pin_memory=True)
            bufs["metric_scaling_factor"] = torch.empty((1, 1), dtype=torch.float32, pin_memory=True)
            if reuse_host_buffers:
                self._host_pool = (key, bufs)
        host = self._host_pool[1] if reuse_host_buffers else bufs
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
        side = self._copy_stream
        main = torch.cuda.current_stream(dev)

        def on_views(outs, v0, v1):
            self._run_postprocess(imgs, outs, H, W, *post_args, post=post, view_lo=v0, view_hi=v1)
            ev = torch.cuda.Event()
            ev.record(main)
            with torch.cuda.stream(side):
                side.wait_event(ev)
                for name, src, k in self._PRED_KEYS:
                    if name == "mask" and not apply_mask:
                        continue
                    host[name][v0:v1].copy_((outs if src == "raw" else post)[k][v0:v1], non_blocking=True)
                if v0 == 0:
                    host["metric_scaling_factor"].copy_(outs["metric_scale"], non_blocking=True)

        with torch.cuda.device(dev):
            side.wait_stream(main)  # a reused pool may still be read by copies of the previous call
            outs = self._run_forward(imgs, H, W, view_lo=view_lo, total_views=total_views, geo=geo,
                                     on_views=on_views, stream_chunk=stream_chunk)
            side.synchronize()
        # the device tensors were only consumed on `side`: tell the caching allocator
        for t in list(outs.values()) + list(post.values()):
            t.record_stream(side)
        preds = []
        for i in range(V):
            d = {name: host[name][i:i + 1] for name, _, _ in self._PRED_KEYS if name != "mask" or apply_mask}
            d["metric_scaling_factor"] = host["metric_scaling_factor"]
            preds.append(d)
        return preds

    # ------------------------------------------------------------------ public API
    @torch.inference_mode()
    def forward(self, views, memory_efficient_inference: bool = False):
        """Raw (un-post-processed) predictions, model.py:1657-2152.  `memory_efficient_inference`
        only changes the dense-head chunking in the reference; outputs are identical, so it is a
        no-op here (the dense head always streams `head_chunk` views through a fixed workspace)."""
        if not self._finalized:
            raise RuntimeError("load_state_dict() / from_pretrained() must be called before inference")
        B = int(views[0]["img"].shape[0])
        if B > 1:  # B independent scenes (see infer)
            per_scene = []
            if self._cfg.aat_nonref_view_pe:  # the reference draws the view-PE rows once per call
                self._pinned_view_pe_rows = self._draw_view_pe_rows(len(views))
            try:
                for b in range(B):
                    scene = [{k: (t[b:b + 1] if torch.is_tensor(t) and t.dim() > 0 and t.shape[0] == B else
                                  (tuple(c[b:b + 1] for c in t) if isinstance(t, tuple) else t))
                              for k, t in v.items()} for v in views]
                    per_scene.append(self.forward(scene, memory_efficient_inference))
            finally:
                self._pinned_view_pe_rows = None
            return [{k: torch.cat([per_scene[b][i][k] for b in range(B)], 0) for k in per_scene[0][i]}
                    for i in range(len(views))]
        imgs, H, W = self._pack_images(views)
        # the reference's forward() draws the per-sample input masks from geometric_input_config's
        # probabilities (model.py:1305-1373); here every geometric input that is present is used, which is
        # what infer() configures (model.py:2170-2197) -- training-time input dropout is out of scope
        geo = self._pack_geometry(views, dict(_ALL_INPUTS), H, W, default_metric=False)
        outs = self._run_forward(imgs, H, W, geo=geo)
        return self._raw_dicts(outs, len(views))

    @torch.inference_mode()
    def infer(self, views: List[Dict[str, Any]], memory_efficient_inference: bool = False,
              use_amp: bool = True, amp_dtype: str = "bf16", apply_mask: bool = True,
              mask_edges: bool = True, edge_normal_threshold: float = 5.0,
              edge_depth_threshold: float = 0.03, apply_confidence_mask: bool = False,
              confidence_percentile: float = 10, ignore_calibration_inputs: bool = False,
              ignore_depth_inputs: bool = False, ignore_pose_inputs: bool = False,
              ignore_depth_scale_inputs: bool = False, ignore_pose_scale_inputs: bool = False,
              view_range=None, total_views: Optional[int] = None, host_outputs: bool = False,
              stream_chunk: int = 4, reuse_host_buffers: bool = False) -> List[Dict[str, torch.Tensor]]:
        """model.py:2206-2355.  The kernels compute with the precision regions of the reference's
        DEFAULT call -- `use_amp=True, amp_dtype="bf16"` (model.py:2287-2302): bf16 operands with fp32
        accumulation and fp32 residual streams in the encoder and the info-sharing transformer, fp32/TF32
        elsewhere.  There is no fp32 trunk and no fp16 path here, so `use_amp=False` and any other
        `amp_dtype` raise NotImplementedError instead of silently returning bf16 results.

        Extension for view-sharded scenes (one process per GPU, torch.distributed initialised):
        `views` are THIS rank's views [view_range[0], view_range[1]) of a scene of `total_views`
        views (parallel.shard_views gives the table); the returned list covers the same views.

        Extension `host_outputs=True`: the prediction tensors are returned in PINNED HOST memory.
        The heads then run `stream_chunk` views at a time and every finished range is post-processed
        and copied out on a second stream while the heads of the later views compute, so the
        107 MB of predictions per 8 views leave the device under the dense head instead of after
        it.  `stream_chunk` may also be a sequence of range sizes (last one repeating); measured at
        8 views (profiles/e2e_variants.py): 4|4 20.28 ms, 4|3|1 20.40, 4|2|2 20.70, 8 20.87, 2|2|2|2
        21.44 -- shrinking the last range shortens the copy-out tail by less than small ranges cost
        the dense head.  The returned tensors are fresh and caller-owned, like the reference's;
        `reuse_host_buffers=True` returns views into ONE pool the model keeps instead (valid until the
        next such call on this model), which saves the pinned allocation of a first call."""
        if not self._finalized:
            raise RuntimeError("load_state_dict() / from_pretrained() must be called before inference")
        if not use_amp or amp_dtype != "bf16":
            raise NotImplementedError(
                f"infer(use_amp={use_amp!r}, amp_dtype={amp_dtype!r}): mapanything_b200 implements the reference's "
                "default mixed-precision path only (use_amp=True, amp_dtype='bf16', model.py:2287-2302); an fp32 or "
                "fp16 run has to use the reference")
        views = validate_input_views_for_inference(views)
        for v in views:
            if v["data_norm_type"][0] != self.data_norm_type:  # encoders/base.py:83-96
                raise AssertionError(f"data_norm_type {v['data_norm_type']} != encoder's {self.data_norm_type}")
        B = int(views[0]["img"].shape[0])
        if B > 1:
            # A per-view batch is B independent scenes (sample b of every view; shown on the live
            # reference by oracle/check_batch_independence.py): run them one after the other and
            # concatenate every prediction along dim 0, which is what the reference returns.
            kw = dict(memory_efficient_inference=memory_efficient_inference, use_amp=use_amp, amp_dtype=amp_dtype,
                      apply_mask=apply_mask, mask_edges=mask_edges, edge_normal_threshold=edge_normal_threshold,
                      edge_depth_threshold=edge_depth_threshold, apply_confidence_mask=apply_confidence_mask,
                      confidence_percentile=confidence_percentile, ignore_calibration_inputs=ignore_calibration_inputs,
                      ignore_depth_inputs=ignore_depth_inputs, ignore_pose_inputs=ignore_pose_inputs,
                      ignore_depth_scale_inputs=ignore_depth_scale_inputs,
                      ignore_pose_scale_inputs=ignore_pose_scale_inputs, view_range=view_range,
                      total_views=total_views, host_outputs=host_outputs, stream_chunk=stream_chunk,
                      reuse_host_buffers=reuse_host_buffers)
            per_scene = []
            if self._cfg.aat_nonref_view_pe:  # the reference draws the view-PE rows once per call
                total = int(total_views or len(views))
                self._pinned_view_pe_rows = self._draw_view_pe_rows(total, sharded=total > len(views))
            try:
                for b in range(B):
                    scene = [{k: (t[b:b + 1] if torch.is_tensor(t) and t.dim() > 0 and t.shape[0] == B else
                                  (tuple(c[b:b + 1] for c in t) if isinstance(t, tuple) else t))
                              for k, t in v.items()} for v in views]
                    preds = self.infer(scene, **kw)
                    # a reused pinned pool is overwritten by the next scene: detach this scene's results
                    per_scene.append([{k: t.clone() for k, t in p.items()} for p in preds]
                                     if host_outputs and reuse_host_buffers else preds)
            finally:
                self._pinned_view_pe_rows = None
            for v in views:  # model.py:2316-2320: the caller's tensors end up on the model device
                for name in list(v.keys()):
                    if name not in _PASS_THROUGH and torch.is_tensor(v[name]):
                        v[name] = v[name].to(self._device, non_blocking=True)
            return [{k: torch.cat([per_scene[b][i][k] for b in range(B)], 0) for k in per_scene[0][i]}
                    for i in range(len(views))]
        flags = dict(use_calibration=not ignore_calibration_inputs, use_depth=not ignore_depth_inputs,
                     use_pose=not ignore_pose_inputs, use_depth_scale=not ignore_depth_scale_inputs,
                     use_pose_scale=not ignore_pose_scale_inputs)  # model.py:2154-2197
        # move the caller's tensors to the model device in place, like model.py:2316-2320; the images
        # go host -> device stack in ONE copy and the caller's dict receives a view of that stack
        V = len(views)
        imgs, H, W = self._pack_images(views)
        for i, v in enumerate(views):
            for name in list(v.keys()):
                if name == "img":
                    v[name] = imgs[i:i + 1]
                elif name not in _PASS_THROUGH and torch.is_tensor(v[name]):
                    v[name] = v[name].to(self._device, non_blocking=True)
        lo = view_range[0] if view_range else 0
        geo = self._pack_geometry(views, flags, H, W, view_lo=lo, total_views=total_views)
        post_args = (apply_mask, mask_edges, edge_normal_threshold, edge_depth_threshold, apply_confidence_mask,
                     confidence_percentile)
        if host_outputs:
            return self._infer_to_host(imgs, H, W, lo, total_views, geo, post_args, stream_chunk,
                                       reuse_host_buffers)
        outs = self._run_forward(imgs, H, W, view_lo=lo, total_views=total_views, geo=geo)
        post = self._run_postprocess(imgs, outs, H, W, *post_args)
        preds = []
        scale = outs["metric_scale"]
        for i in range(V):
            d = {
                "pts3d": outs["pts3d"][i:i + 1],
                "pts3d_cam": outs["pts3d_cam"][i:i + 1],
                "ray_directions": outs["ray_directions"][i:i + 1],
                "depth_along_ray": outs["depth_along_ray"][i:i + 1],
                "cam_trans": post["cam_trans"][i:i + 1],
                "cam_quats": post["cam_quats"][i:i + 1],
                "metric_scaling_factor": scale,
                "conf": outs["conf"][i:i + 1],
                "non_ambiguous_mask": outs["non_ambiguous_mask"][i:i + 1],
                "non_ambiguous_mask_logits": outs["mask_logits"][i:i + 1],
                "img_no_norm": post["img_no_norm"][i:i + 1],
                "depth_z": post["depth_z"][i:i + 1],
                "intrinsics": post["intrinsics"][i:i + 1],
                "camera_poses": post["camera_poses"][i:i + 1],
            }
            if apply_mask:
                d["mask"] = post["mask"][i:i + 1]
            preds.append(d)
        return preds
