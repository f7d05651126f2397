"""TEST INFRASTRUCTURE ONLY -- live-reference shim (works only where /root/reference exists).

Imports the *unmodified* reference `MapAnything` class from /root/reference so that
`oracle/make_golden.py` can generate the golden vectors under tests/golden/ and so that the
restatement in `oracle/mapanything_oracle.py` can be pinned against the real thing.

Nothing in the product package, the `-m gpu` tests, `smoke()` or `bench.py` imports this file:
/root/reference does not exist on the GPU box.

Recipe:
  1. stub the five viz/config modules the reference imports but never executes on the inference
     path (omegaconf, rerun, trimesh, viser, matplotlib);
  2. redirect `torch.hub.load("facebookresearch/dinov2", "dinov2_vitl14")`
     (uniception/models/encoders/dinov2.py:85-91) to the vendored, architecture-identical
     `mapanything/models/external/dinov2/hub/backbones.py:dinov2_vitl14(pretrained=False)`;
  3. build `MapAnything(**cfg)` from a literal dict mirroring the resolved Hydra config
     (configs/model/mapanything.yaml and friends).
"""
import copy
import importlib.abc
import importlib.machinery
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("MAB_REFERENCE_ROOT", "/root/reference")
_STUBBED = {"omegaconf", "rerun", "trimesh", "viser", "matplotlib"}
INF = float("inf")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "mapanything"))


class _StubModule(types.ModuleType):
    __path__ = []  # behave as a package so that `import a.b as c` works

    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        full = f"{self.__name__}.{name}"
        mod = sys.modules.get(full)
        if mod is None:
            mod = _StubModule(full)
            sys.modules[full] = mod
        return mod

    def __call__(self, *a, **k):
        return self


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, name, path=None, target=None):
        if name.split(".")[0] in _STUBBED:
            return importlib.machinery.ModuleSpec(name, self, is_package=True)
        return None

    def create_module(self, spec):
        return _StubModule(spec.name)

    def exec_module(self, module):
        pass


_installed = False


def install():
    """Install import stubs + hub shim. Idempotent."""
    global _installed
    if _installed:
        return
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    import torch  # noqa: F401  (must be imported before the finder is installed)
    import torchvision  # noqa: F401

    sys.meta_path.append(_StubFinder())
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)

    from mapanything.models.external.dinov2.hub.backbones import dinov2_vitl14

    def _hub_load(repo, name, *a, **k):
        assert name == "dinov2_vitl14", name
        return dinov2_vitl14(pretrained=False)

    torch.hub.load = _hub_load
    _installed = True


def reference_config():
    """The production configuration."""
    cfg = dict(
        name="mapanything",
        encoder_config=dict(
            encoder_str="dinov2", name="dinov2_large", data_norm_type="dinov2", size="large",
            with_registers=False, uses_torch_hub=True, gradient_checkpointing=False,
        ),
        info_sharing_config=dict(
            model_type="alternating_attention", model_return_type="intermediate_features",
            custom_positional_encoding=None,
            module_args=dict(
                name="aat_24_layers_ifr", indices=[11, 17], norm_intermediate=True,
                size="24_layers", depth=24, distinguish_ref_and_non_ref_views=True,
                gradient_checkpointing=False,
            ),
        ),
        pred_head_config=dict(
            type="dpt+pose",
            feature_head=dict(feature_dim=256, hooks=[0, 1, 2, 3], checkpoint_gradient=False),
            regressor_head=dict(input_feature_dim=256, output_dim=6, checkpoint_gradient=False),
            pose_head=dict(num_resconv_block=2, rot_representation_dim=4),
            scale_head=dict(output_dim=1),
            adaptor_type="raydirs+depth+pose+confidence+mask",
            dpt_adaptor=dict(
                name="dpt_adaptor", ray_directions_mode="linear",
                ray_directions_normalize_to_unit_sphere=True,
                ray_directions_normalize_to_unit_image_plane=False,
                ray_directions_vmin=-INF, ray_directions_vmax=INF,
                ray_directions_clamp_min_of_z_dir=False, ray_directions_z_dir_min=-INF,
                depth_mode="exp", depth_vmin=0, depth_vmax=INF,
                confidence_type="exp", confidence_vmin=1, confidence_vmax=INF,
            ),
            pose_adaptor=dict(
                name="pose_adaptor", cam_trans_mode="linear", cam_trans_vmin=-INF,
                cam_trans_vmax=INF, quaternions_mode="linear", quaternions_normalize=True,
                quaternions_vmin=-INF, quaternions_vmax=INF,
            ),
            scale_adaptor=dict(name="scale_adaptor", mode="exp", vmin=1e-8, vmax=INF),
            gradient_checkpointing=False,
        ),
        geometric_input_config=dict(
            ray_dirs_encoder_config=dict(name="ray_dirs_encoder", in_chans=3,
                                         encoder_str="dense_rep_encoder", apply_pe=False),
            depth_encoder_config=dict(name="depth_encoder", in_chans=1,
                                      encoder_str="dense_rep_encoder", apply_pe=False),
            cam_rot_encoder_config=dict(name="cam_rot_quats_encoder", in_chans=4,
                                        encoder_str="global_rep_encoder"),
            cam_trans_encoder_config=dict(name="cam_trans_encoder", in_chans=3,
                                          encoder_str="global_rep_encoder"),
            scale_encoder_config=dict(name="scale_encoder", in_chans=1,
                                      encoder_str="global_rep_encoder"),
            overall_prob=0, dropout_prob=1, ray_dirs_prob=0, depth_prob=0, cam_prob=0,
            sparse_depth_prob=0, sparsification_removal_percent=0,
            depth_scale_norm_all_prob=0, pose_scale_norm_all_prob=0,
        ),
    )
    return cfg


def build_reference_model(state_dict=None, module_args=None, model_type=None):
    """Construct the reference MapAnything (CPU, fp32, eval); optionally load a state dict.
    module_args: overrides of info_sharing_config.module_args (the configs/model/info_sharing/*.yaml
    variants, e.g. use_entropy_scaling=True, distinguish_ref_and_non_ref_views=False)."""
    install()
    from mapanything.models.mapanything.model import MapAnything

    cfg = copy.deepcopy(reference_config())
    if model_type is not None and model_type != cfg["info_sharing_config"]["model_type"]:
        # another transformer class (gat_ifr_24_layers.yaml): its module_args replace the AAT ones
        cfg["info_sharing_config"]["model_type"] = model_type
        cfg["info_sharing_config"]["module_args"] = dict(module_args or {})
    else:
        cfg["info_sharing_config"]["module_args"].update(module_args or {})
    model = MapAnything(**cfg).eval()
    if state_dict is not None:
        sd = dict(state_dict)
        if not cfg["info_sharing_config"]["module_args"].get("distinguish_ref_and_non_ref_views", True):
            sd.pop("info_sharing.view_pos_table", None)  # the buffer only exists with the reference-view PE
        missing, unexpected = model.load_state_dict(sd, strict=True)
        assert not missing and not unexpected
    return model
