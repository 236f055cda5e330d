"""CPU: the CycleGAN oracle restatement itself (oracle/cyclegan_oracle.py; cyclegan/CycleGAN.py:37-208,325-382).  The
reference has no tests and TensorFlow / tensorflow_addons cannot run here, so parity with them is UNPINNED; these tests pin
the restatement's own definitions: layer arithmetic against first principles, variable inventory, step bookkeeping."""
import torch
import torch.nn.functional as F

from oracle import cyclegan_oracle as CO


def test_row_norm_is_per_image_row():
    x = torch.randn(2, 5, 7, 3, dtype=torch.float64)
    g, b = torch.randn(5, dtype=torch.float64), torch.randn(5, dtype=torch.float64)
    y = CO.row_norm(x, g, b)
    for n in range(2):
        for h in range(5):
            r = x[n, h]
            want = (r - r.mean()) / torch.sqrt(r.var(unbiased=False) + 1e-3) * g[h] + b[h]
            assert torch.allclose(y[n, h], want, atol=1e-12)


def test_conv_transpose_k3s2_same_is_the_gradient_of_a_same_conv():
    # Keras Conv2DTranspose(3, strides 2, 'same') = input gradient of Conv2D(3, strides 2, 'same'), whose single padding
    # row / column sits at the END for even sizes
    w = torch.randn(3, 3, 4, 5, dtype=torch.float64)          # as conv kernel [kh,kw,Cin=4,Cout=5] = convT [kh,kw,Cout=4,Cin=5]
    xx = torch.randn(1, 6, 6, 5, dtype=torch.float64)
    inp = torch.zeros(1, 12, 12, 4, dtype=torch.float64, requires_grad=True)
    pad = F.pad(inp.permute(0, 3, 1, 2), (0, 1, 0, 1))
    out = F.conv2d(pad, w.permute(3, 2, 0, 1), stride=2)
    out.backward(xx.permute(0, 3, 1, 2))
    got = CO.conv2d_transpose_k3s2_same(xx, w, None)
    assert got.shape == (1, 12, 12, 4) and float((got - inp.grad).abs().max()) < 1e-12


def test_reflect_pad_and_shapes():
    x = torch.arange(16.0).reshape(1, 4, 4, 1)
    p = CO.reflect_pad(x)
    assert p.shape == (1, 6, 6, 1) and float(p[0, 0, 0, 0]) == float(x[0, 1, 1, 0]) and float(p[0, 5, 2, 0]) == float(x[0, 2, 1, 0])
    assert CO.disc_sizes(128) == [128, 63, 30, 14, 6, 3]          # PatchGAN logits 3 x 3 at the reference's 128 x 128
    gs = CO.generator_var_shapes()
    ds = CO.discriminator_var_shapes()
    assert len(gs) == 96 and len(ds) == 16
    n = lambda shapes: sum(int(torch.tensor(s).prod()) for _, s in shapes)
    assert n(gs) == 7 * 7 * 3 * 64 + 64 + 3 * 3 * 64 * 128 + 128 + 3 * 3 * 128 * 256 + 256 + 18 * (9 * 256 * 256 + 256) \
        + 9 * 256 * 128 + 128 + 9 * 128 * 64 + 64 + 49 * 64 * 3 + 3 + 2 * (128 + 64 + 32 + 18 * 32 + 64 + 128 + 128)


def test_train_step_losses_and_bookkeeping():
    o = CO.OracleCycleGAN((96, 96, 3), seed=3)
    g = torch.Generator().manual_seed(0)
    x = torch.rand((1, 96, 96, 3), generator=g) * 2 - 1
    y = torch.rand((1, 96, 96, 3), generator=g) * 2 - 1
    w0 = [v.clone() for v in o.g]
    col = {}
    r = o.train_step(x, y, col)
    assert abs(r["total_gen_g_loss"] - (r["gen_g_loss"] + r["total_cycle_loss"] + r["identity_loss_g"])) < 1e-5
    assert abs(r["total_gen_f_loss"] - (r["gen_f_loss"] + r["total_cycle_loss"] + r["identity_loss_f"])) < 1e-5
    assert 0.3 < r["gen_g_loss"] < 1.5 and r["total_cycle_loss"] > 0
    assert all(o.opt[k].t == 1 for k in ("g", "f", "dx", "dy"))
    # first Adam step with beta_1 = 0.5: |dw| = lr wherever the gradient is clearly non-zero
    gk = col["grads_g"][0]
    moved = (o.g[0] - w0[0]).abs()
    assert torch.allclose(moved[gk.abs() > 1e-5], torch.full_like(moved[gk.abs() > 1e-5], 2e-4), rtol=1e-2)
