"""The reference has no test suite; its only tests are the `if __name__ == "__main__"` self-checks of
the modules on the hot path.  They are re-stated here on the oracle (CPU), so the
structural facts the CUDA path's data layout relies on stay asserted:

* alternating_attention_transformer.py:781-818  `(B, V*T, C).reshape(B*V, T, C)` keeps (view, token) order
  -- the token-major layout "row v*T + t" of DESIGN.md section 3;
* alternating_attention_transformer.py:931-943  with norm_intermediate, a tap taken at the last layer equals
  the final output -- taps and final output share ONE LayerNorm (`info_sharing.norm`);
* pose_head.py:176-179 (B,7), mlp_head.py:109-112 scalar per scene, dense_rep_encoder.py:306-313 1024
  channels on the patch grid, global_rep_encoder.py:108-115 (B,1024)."""
import torch

from oracle import mapanything_oracle as orc


def test_reshape_for_frame_attention_keeps_view_token_order():
    B, V, T, Cc = 1, 3, 5, 4
    x = torch.zeros(B, V * T, Cc)
    for v in range(V):
        for t in range(T):
            x[0, v * T + t] = 100 * v + t
    y = x.reshape(B * V, T, Cc)  # what the odd (frame) layers attend over
    for v in range(V):
        assert torch.equal(y[v, :, 0], 100 * v + torch.arange(T, dtype=torch.float32))
    assert torch.equal(y.reshape(B, V * T, Cc), x)


def test_last_normed_intermediate_equals_final(synthetic_sd):
    torch.manual_seed(0)
    feats = torch.randn(2, 6, 1024)  # 2 views of 2 x 3 patches
    with torch.inference_mode():
        final, inter, scale_tok = orc.alternating_attention_transformer(feats, synthetic_sd, indices=(11, 23))
    assert final.shape == (2, 6, 768) and scale_tok.shape == (768,)
    assert len(inter) == 2 and inter[0].shape == (2, 6, 768)
    assert torch.equal(inter[1], final)
    assert not torch.equal(inter[0], final)


def test_head_and_encoder_output_shapes(synthetic_sd):
    torch.manual_seed(1)
    with torch.inference_mode():
        t, q = orc.pose_head(torch.randn(3, 768, 2, 3), synthetic_sd)
        assert t.shape == (3, 3) and q.shape == (3, 4)
        assert torch.allclose(q.norm(dim=1), torch.ones(3), atol=1e-6)      # unit quaternions
        s = orc.scale_head(torch.randn(768), synthetic_sd)
        assert s.shape == (1, 1) and float(s) >= 1e-8
        d = orc.dense_rep_encoder(torch.randn(2, 3, 28, 42), synthetic_sd, "ray_dirs_encoder")
        assert d.shape == (2, 2 * 3, 1024)
        d1 = orc.dense_rep_encoder(torch.randn(2, 1, 28, 42), synthetic_sd, "depth_encoder")
        assert d1.shape == (2, 6, 1024)
        for prefix, cin in (("cam_rot_encoder", 4), ("cam_trans_encoder", 3), ("depth_scale_encoder", 1),
                            ("cam_trans_scale_encoder", 1)):
            assert orc.global_rep_encoder(torch.randn(5, cin), synthetic_sd, prefix).shape == (5, 1024)
