"""CPU tests of the WGAN oracle (oracle/wgan_oracle.py; reference wasserstein_gan/WGAN.py): Keras semantics restated
by hand, and the frozen fixture tests/golden/wgan_oracle_tiny.npz."""
import importlib.util
import os

import numpy as np
import torch
import torch.nn.functional as F

from oracle import wgan_oracle as W

HERE = os.path.dirname(os.path.abspath(__file__))


def test_wgan_golden_fixture_reproduces():
    spec = importlib.util.spec_from_file_location("make_wgan_golden", os.path.join(HERE, "golden", "make_wgan_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    got = mod.run()
    ref = np.load(os.path.join(HERE, "golden", "wgan_oracle_tiny.npz"))
    for k in ref.files:
        assert np.allclose(got[k], ref[k], rtol=1e-6, atol=1e-8), k


def test_variable_order_and_counts():
    cs = W.critic_var_shapes((144, 256, 3))
    assert [n for n, _ in cs[:4]] == ["conv1/kernel", "conv1/bias", "bn1/gamma", "bn1/beta"]
    assert cs[28] == ("dense/kernel", (294912, 1)) and len(cs) == 30
    gs = W.generator_var_shapes((144, 256, 3))
    assert gs[0] == ("dense/kernel", (128, 294912)) and gs[-1] == ("conv4/kernel", (3, 3, 64, 3)) and len(gs) == 11
    assert len(W.OracleWGAN.critic_bn_indices()) == 14


def test_batchnorm_training_matches_torch_and_moves_statistics():
    g = torch.Generator().manual_seed(3)
    x = torch.randn(3, 4, 5, 8, generator=g, dtype=torch.float64) * 2 + 1
    gamma, beta = torch.rand(8, generator=g, dtype=torch.float64) + 0.5, torch.randn(8, generator=g, dtype=torch.float64)
    moving = [torch.zeros(8, dtype=torch.float64), torch.ones(8, dtype=torch.float64)]
    y = W.batchnorm(x, gamma, beta, moving, True)
    rm, rv = torch.zeros(8, dtype=torch.float64), torch.ones(8, dtype=torch.float64)
    # torch: momentum is the weight of the NEW value (1 - Keras momentum); running_var takes the unbiased variance
    yt = F.batch_norm(x.permute(0, 3, 1, 2), rm, rv, gamma, beta, True, 1 - W.BN_MOMENTUM, W.BN_EPS).permute(0, 2, 3, 1)
    assert torch.allclose(y, yt, atol=1e-12)
    assert torch.allclose(moving[0], rm, atol=1e-12) and torch.allclose(moving[1], rv, atol=1e-12)
    yi = W.batchnorm(x, gamma, beta, moving, False)
    assert torch.allclose(yi, gamma * (x - moving[0]) / torch.sqrt(moving[1] + 1e-3) + beta)


def test_rmsprop_by_hand_and_clip():
    w, g, r = torch.tensor([0.5]), torch.tensor([2.0]), torch.zeros(1)
    W.rmsprop_step(w, g, r, lr=5e-5)
    assert abs(float(r) - 0.4) < 1e-7                                  # 0.9*0 + 0.1*4
    assert abs(float(w) - (0.5 - 5e-5 * 2.0 / (0.4 ** 0.5 + 1e-7))) < 1e-7
    W.rmsprop_step(w, g, r, lr=5e-5)
    assert abs(float(r) - 0.76) < 1e-6


def test_train_on_batch_trainable_sets():
    """critic.train_on_batch moves everything and clips only conv kernels; gan.train_on_batch leaves the critic's
    conv / Dense variables alone but moves its BatchNorm gamma/beta and the moving statistics (WGAN.py:140-142)."""
    o = W.OracleWGAN((16, 32, 3), seed=5)
    g = torch.Generator().manual_seed(1)
    x = torch.rand(4, 16, 32, 3, generator=g) * 2 - 1
    before = [v.clone() for v in o.cv]
    o.train_critic_on_batch(x, -1.0)
    assert all(float(o.cv[4 * i].abs().max()) <= W.CLIP + 1e-9 for i in range(7))
    assert not torch.equal(o.cv[28], before[28]) and float(o.cv[28].abs().max()) > W.CLIP   # Dense not clipped
    mid = [v.clone() for v in o.cv]
    mm = o.c_moving[0][0].clone()
    gv0 = [v.clone() for v in o.gv]
    o.train_gan_on_batch(torch.randn(4, 128, generator=g))
    for i in range(30):
        changed = not torch.equal(o.cv[i], mid[i])
        assert changed == (i in o.critic_bn_indices()), i
    assert all(not torch.equal(a, b) for a, b in zip(o.gv, gv0))
    assert not torch.equal(o.c_moving[0][0], mm)
    # predict uses the moving statistics: differs from a training-mode forward
    z = torch.randn(4, 128, generator=g)
    a = o.predict(z)
    b = W.generator_forward(o.gv, [[m[0].clone(), m[1].clone()] for m in o.g_moving], z, True, (16, 32, 3))
    assert float((a - b).abs().max()) > 1e-4


def test_wasserstein_loss():
    y = torch.tensor([[1.0], [3.0]])
    assert float(W.wasserstein_loss(-1.0, y)) == -2.0


def test_fid_oracle_lowrank_identity_and_pooling():
    """tr sqrtm(cov_f cov_r) == nuclear norm of Xc Yc^T / sqrt((n-1)(m-1)) (what the CUDA path evaluates), checked
    against the reference's scipy.linalg.sqrtm arithmetic; AveragePooling2D(8, 8) 'valid' drops trailing rows."""
    import numpy as np
    from oracle import evaluation_oracle as E
    rng = np.random.default_rng(0)
    a, b = rng.normal(size=(40, 24)), rng.normal(size=(50, 24)) * 1.3 + 0.2
    fid = E.calculate_fid_from_features(a, b)
    alt = (np.sum((a.mean(0) - b.mean(0)) ** 2) + np.trace(np.cov(a, rowvar=False)) + np.trace(np.cov(b, rowvar=False))
           - 2 * E.trace_sqrt_product_lowrank(a, b))
    assert abs(fid - alt) < 1e-8 * abs(fid)
    assert abs(E.calculate_fid_from_features(a, a)) < 1e-6
    x = torch.arange(2 * 18 * 32 * 3, dtype=torch.float64).reshape(2, 18, 32, 3)
    p = E.avg_pool_8x8(x)
    assert tuple(p.shape) == (2, 2, 4, 3)
    assert torch.allclose(p[0, 1, 2], x[0, 8:16, 16:24].mean(dim=(0, 1)))
