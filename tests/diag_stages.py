"""TEST INFRASTRUCTURE -- stage-by-stage error attribution on the GPU box.

    python -m tests.diag_stages [num_views H W]

Runs, on the same synthetic weights and views,
  (F) the oracle on the GPU in fp32 (TF32 off)                        = ground truth,
  (A) the oracle on the GPU with the reference's precision semantics  = "reference bf16 path":
      encoder + info-sharing under torch.autocast(bf16), fusion norm / heads / adaptors in fp32
      with TF32 matmul+cuDNN allowed (model.py:89-93, 1377, 1774, 2335-2338),
  (M) the CUDA path (libmab200),
and prints the relative L2 error of (A) and (M) against (F) at every stage tap.  (A) is the yardstick
for what "within bf16 tolerance of the reference's own PyTorch path" means on these weights.
"""
import ctypes as C
import sys

import torch

from mapanything_b200 import MapAnything, _lib
from oracle import mapanything_oracle as orc
from oracle import weights
from tests import helpers


def debug_buffer(model, name):
    ptr, n, dt = C.c_void_p(), C.c_longlong(), C.c_int()
    _lib.check(model._lib.mab_debug_buffer(model._handle, name.encode(), C.byref(ptr), C.byref(n), C.byref(dt)))
    torch.cuda.synchronize()

    class _Dev:  # wrap the raw device pointer through the CUDA array interface
        pass

    o = _Dev()
    if dt.value == 0:
        o.__cuda_array_interface__ = {"shape": (n.value,), "typestr": "<f4", "data": (ptr.value, False), "version": 2}
        out = torch.as_tensor(o, device=model.device).clone()
    else:  # bf16 has no numpy typestr: view as int16 and reinterpret
        o.__cuda_array_interface__ = {"shape": (n.value,), "typestr": "<i2", "data": (ptr.value, False), "version": 2}
        out = torch.as_tensor(o, device=model.device).clone().view(torch.bfloat16)
    return out.float()


def oracle_stages(sd, views, autocast):
    """Returns dict of taps computed with the reference's precision regions."""
    dev = "cuda"
    V = len(views)
    H, W = views[0]["img"].shape[-2:]
    h, w = H // 14, W // 14
    imgs = torch.cat([v["img"] for v in views], 0).to(dev)
    torch.backends.cuda.matmul.allow_tf32 = bool(autocast)
    torch.backends.cudnn.allow_tf32 = bool(autocast)
    taps = {}
    ctx = (lambda: torch.autocast("cuda", dtype=torch.bfloat16)) if autocast else (lambda: torch.autocast("cuda", enabled=False))
    with torch.inference_mode():
        with ctx():
            enc = orc.dinov2_encoder(imgs, sd).float()
        with torch.autocast("cuda", enabled=False):
            fused = orc._ln(enc, sd, "fusion_norm_layer")
        with ctx():
            final, inter, scale_tok = orc.alternating_attention_transformer(fused, sd)
            final, inter, scale_tok = final.float(), [t.float() for t in inter], scale_tok.float()
        with torch.autocast("cuda", enabled=False):
            def nchw(t):
                return t.reshape(V, h, w, -1).permute(0, 3, 1, 2).contiguous()
            layers = [nchw(fused), nchw(inter[0]), nchw(inter[1]), nchw(final)]
            feat8 = orc.dpt_feature(layers, sd)
            logits = orc.dpt_regressor(feat8, sd, (H, W))
            t, q = orc.pose_head(layers[3], sd)
            scale = orc.scale_head(scale_tok, sd)
    taps.update(fused=fused, tap0=inter[0], tap1=inter[1], final=final, scale_token=scale_tok,
                feat8=feat8.permute(0, 2, 3, 1), logits=logits.permute(0, 2, 3, 1), pose_t=t, pose_q=q,
                scale=scale)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = True
    return taps


def main():
    n, H, W = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (3, 70, 98)
    sd_cpu = weights.synthetic_state_dict()
    sd = {k: v.cuda() for k, v in sd_cpu.items()}
    views = weights.synthetic_views(n, H, W, seed=1)
    F32 = oracle_stages(sd, views, autocast=False)
    AMP = oracle_stages(sd, views, autocast=True)
    model = MapAnything(device="cuda:0", head_chunk=n)
    model.load_state_dict(sd_cpu)
    imgs, _, _ = model._pack_images([dict(v) for v in views])
    outs = model._run_forward(imgs, H, W, want_logits=True)
    torch.cuda.synchronize()
    M = {k: debug_buffer(model, k) for k in ("fused", "tap0", "tap1", "final", "scale_token", "feat8")}
    M["logits"] = outs["logits6"]
    M["pose_t"] = outs["pose"][:, 0:3]
    M["pose_q"] = outs["pose"][:, 3:7]
    M["scale"] = outs["metric_scale"]
    print(f"views={n} {H}x{W}   rel-L2 vs fp32 oracle:   reference-precision(A)    libmab200(M)")
    for k in ("fused", "tap0", "tap1", "final", "scale_token", "feat8", "logits", "pose_t", "pose_q", "scale"):
        f = F32[k].float().reshape(-1)
        a = AMP[k].float().reshape(-1)
        m = M[k].float().reshape(-1)[: f.numel()]
        print(f"  {k:12s}  A {helpers.rel_l2(a.cpu(), f.cpu()):.3e}   M {helpers.rel_l2(m.cpu(), f.cpu()):.3e}")
    for name, t in (("A", AMP), ("M", M)):
        ang = helpers.quat_angle_deg(t["pose_q"].cpu().numpy(), F32["pose_q"].cpu().numpy())
        print(f"  pose angle vs fp32: {name} {ang:.4f} deg")
    lf = F32["logits"].float()
    for name, t in (("A", AMP), ("M", M)):
        l = t["logits"].float().reshape(lf.shape)
        per = [(helpers.rel_l2(l[..., c].cpu(), lf[..., c].cpu())) for c in range(6)]
        print(f"  logits per channel {name}: " + " ".join(f"{e:.2e}" for e in per))


if __name__ == "__main__":
    main()
