"""imagegeneration_b200 — the SNDCGAN training step of jonathan-schilling/imageGeneration on B200 (sm_100a).

Public surface (mirrors sndcgan/SNDCGAN.py, sndcgan/Trainer.py, sndcgan/generator_output.py of the reference):
    make_dcgan_generator, make_dcgan_discriminator, get_dataset, SNDCGAN (train_step / train), create_samples, Adam,
    BinaryCrossentropy (the `loss_fn`), save_model / load_model (the per-checkpoint model exports)
The CUDA library (libb200gan.so, C ABI in include/b200gan.h) is loaded on first use; there is no CPU fallback.
"""
__all__ = ["make_dcgan_generator", "make_dcgan_discriminator", "get_dataset", "SNDCGAN", "SNDCGANEngine",
           "create_samples", "Adam", "BinaryCrossentropy", "save_model", "load_model"]


def __getattr__(name):
    if name in ("make_dcgan_generator", "make_dcgan_discriminator", "SNDCGANEngine", "create_samples", "Adam"):
        from . import sndcgan
        return getattr(sndcgan, name)
    if name in ("SNDCGAN", "get_dataset", "save_model", "load_model"):
        from . import trainer
        return getattr(trainer, name)
    if name == "BinaryCrossentropy":
        from .torch_ops import BinaryCrossentropy
        return BinaryCrossentropy
    raise AttributeError(name)
