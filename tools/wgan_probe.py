"""Prints the teacher-forced loss differences engine vs oracle (fp32 and precision-emulating) — threshold setting."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import test_wgan_gpu as T

for emu in (False, True):
    eng, orc = T._make_pair(emu=emu)
    g = torch.Generator().manual_seed(123)
    reals = [torch.rand((T.BATCH, *T.IMAGE), generator=g) * 2 - 1 for _ in range(2)]
    zs = [torch.randn((T.BATCH, 128), generator=g) for _ in range(3)]
    for it in range(2):
        for x, z in zip(reals, zs[:2]):
            T._sync(eng, orc)
            a, b = eng.train_critic_on_batch(x.cuda(), -1.0), orc.train_critic_on_batch(x, -1.0)
            sc = float(eng.critic_model.logits.abs().mean())
            print(emu, it, "real", float(a), float(b), "rel", abs(float(a) - float(b)) / sc, "logit std", float(eng.critic_model.logits.std()))
            T._sync(eng, orc)
            fo = orc.predict(z)
            a, b = eng.train_critic_on_batch(fo.cuda(), 1.0), orc.train_critic_on_batch(fo, 1.0)
            sc = float(eng.critic_model.logits.abs().mean())
            print(emu, it, "fake", float(a), float(b), "rel", abs(float(a) - float(b)) / sc, "img std over batch", float(fo.std(0).mean()), "img std", float(fo.std()))
        T._sync(eng, orc)
        a, b = eng.train_gan_on_batch(zs[2].cuda()), orc.train_gan_on_batch(zs[2])
        sc = float(eng.critic_model.logits.abs().mean())
        print(emu, it, "gan", float(a), float(b), "rel", abs(float(a) - float(b)) / sc)
