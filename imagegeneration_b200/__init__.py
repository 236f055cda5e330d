"""imagegeneration_b200 — the SNDCGAN training step of jonathan-schilling/imageGeneration on B200 (sm_100a).

Public surface (mirrors sndcgan/SNDCGAN.py, sndcgan/Trainer.py, sndcgan/generator_output.py of the reference):
    make_dcgan_generator, make_dcgan_discriminator, SNDCGAN (train_step / train), create_samples, Adam
The CUDA library (libb200gan.so, C ABI in include/b200gan.h) is loaded on first use; there is no CPU fallback.
"""
__all__ = ["make_dcgan_generator", "make_dcgan_discriminator", "SNDCGAN", "SNDCGANEngine", "create_samples", "Adam"]


def __getattr__(name):
    if name in ("make_dcgan_generator", "make_dcgan_discriminator", "SNDCGANEngine", "create_samples", "Adam"):
        from . import sndcgan
        return getattr(sndcgan, name)
    if name == "SNDCGAN":
        from .trainer import SNDCGAN
        return SNDCGAN
    raise AttributeError(name)
