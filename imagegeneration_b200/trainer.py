"""Drop-in for the reference's `sndcgan/Trainer.py` CLI and the non-kernel parts of its `SNDCGAN` class
(sndcgan/SNDCGAN.py:131-212, :272-335): dataset iterator, output directory handling, losses pickle, checkpoints,
epoch loop.  The training step itself is SNDCGANEngine.train_step (imagegeneration_b200/sndcgan.py).

    python -m imagegeneration_b200.trainer <bSize> <epochs> [-cf N] [-d DIR] [-x DATA] [-r RATE] [-ld LR] [-lg LR]
                                                           [-lo NAME] [-ct]   (+ opt-in extras, defaults = reference)
"""
from __future__ import annotations

import argparse
import glob
import os
import pickle
import shutil
from time import gmtime, strftime, time

import numpy as np
import torch

from .sndcgan import SNDCGANEngine, create_samples

img_height = 144          # sndcgan/Trainer.py:5-8
img_width = 256
image_size = (img_height, img_width, 3)
z_size = 128


# ----------------------------------------------------------------------------------------------------------
# dataset  (SNDCGAN.py:131-145)
# ----------------------------------------------------------------------------------------------------------
class ImageFolderDataset:
    """`image_dataset_from_directory(seed=123, crop_to_aspect_ratio=True)` + `Rescaling(1/127.5, offset=-1)` +
    cache/shuffle: decodes every image once (centre crop to the target aspect ratio, bilinear resize), keeps the
    uint8 cache in memory and yields shuffled fp32 NHWC batches in [-1, 1] with dummy labels."""

    def __init__(self, root, batch_size, size, seed=123):
        from PIL import Image
        exts = ("*.jpg", "*.jpeg", "*.png", "*.bmp")
        files = sorted(f for e in exts for f in glob.glob(os.path.join(root, "**", e), recursive=True))
        if not files:
            raise FileNotFoundError(f"no images under {root!r} (use --synthetic for generated data)")
        h, w = size[:2]
        imgs = np.empty((len(files), h, w, 3), dtype=np.uint8)
        for i, f in enumerate(files):
            im = Image.open(f).convert("RGB")
            iw, ih = im.size
            if iw * h > ih * w:      # too wide: crop width
                cw = int(ih * w / h)
                x0 = (iw - cw) // 2
                im = im.crop((x0, 0, x0 + cw, ih))
            else:
                ch = int(iw * h / w)
                y0 = (ih - ch) // 2
                im = im.crop((0, y0, iw, y0 + ch))
            imgs[i] = np.asarray(im.resize((w, h), Image.BILINEAR))
        self.images = torch.from_numpy(imgs)
        self.batch_size = batch_size
        self.gen = torch.Generator().manual_seed(seed)

    def __len__(self):
        return (len(self.images) + self.batch_size - 1) // self.batch_size

    def __iter__(self):
        perm = torch.randperm(len(self.images), generator=self.gen)
        for i in range(0, len(perm), self.batch_size):
            idx = perm[i:i + self.batch_size]
            x = self.images[idx].to(torch.float32) * (1.0 / 127.5) - 1.0
            yield x.pin_memory() if torch.cuda.is_available() else x, torch.zeros(len(idx), dtype=torch.int64)


class SyntheticDataset:
    """U(-1,1) images of the reference's training shape (Open Images is not available offline); 7 244 images per
    epoch by default = the reference's usable set (dataset_creator/results_open_images/stats.json)."""

    def __init__(self, batch_size, size, n_images=7244, seed=123):
        self.batch_size, self.size, self.n = batch_size, size, n_images
        self.gen = torch.Generator().manual_seed(seed)

    def __len__(self):
        return self.n // self.batch_size

    def __iter__(self):
        for _ in range(len(self)):
            x = torch.rand((self.batch_size, *self.size), generator=self.gen) * 2 - 1
            yield x.pin_memory() if torch.cuda.is_available() else x, torch.zeros(self.batch_size, dtype=torch.int64)


def get_dataset(dataset, batch_size, image_size):
    """SNDCGAN.py:131: directory of images -> iterable of (fp32 NHWC in [-1,1], labels); "synthetic" -> generated."""
    if dataset == "synthetic" or str(dataset).startswith("synthetic:"):
        n = int(str(dataset).split(":")[1]) if ":" in str(dataset) else 7244
        return SyntheticDataset(batch_size, tuple(image_size), n)
    return ImageFolderDataset(dataset, batch_size, tuple(image_size))


# ----------------------------------------------------------------------------------------------------------
# SNDCGAN  (SNDCGAN.py:148-335)
# ----------------------------------------------------------------------------------------------------------
class SNDCGAN(SNDCGANEngine):
    def __init__(self, dir_path, dataset, batch_size, dropout, learning_rate_disc, learning_rate_gen, live_output,
                 continue_, image_size, z_size, **engine_kwargs):
        if not continue_ and os.path.exists(dir_path):          # :152-153 (yes, it deletes the directory)
            shutil.rmtree(dir_path)
        if not os.path.exists(dir_path):
            os.mkdir(dir_path)
        self.dir_path = dir_path
        self.train_ds = get_dataset(dataset, batch_size, image_size)
        super().__init__(batch_size, dropout=dropout, learning_rate_disc=learning_rate_disc,
                         learning_rate_gen=learning_rate_gen, image_size=image_size, z_size=z_size, **engine_kwargs)
        self.live_preview_file = os.path.join(live_output + ".png")
        self.losses_file = os.path.join(dir_path, "losses.pickle")
        if os.path.exists(self.losses_file):
            with open(self.losses_file, mode="rb") as f:
                self.losses = pickle.load(f)
        else:
            self.losses = {"epoch": [], "avg_g_loss": [], "avg_d_loss": [], "d_real": [], "d_fake": []}
        self.checkpoint_path = os.path.join(dir_path, "checkpoints")
        os.makedirs(self.checkpoint_path, exist_ok=True)
        latest = self._latest_checkpoint()
        if continue_ and latest:                                 # :192-196
            self.start_epoch = int(os.path.basename(latest).split("-")[-1].split(".")[0]) + 1
            self._restore(latest)
            print("Latest checkpoint restored!!")
        else:
            self.start_epoch = 0
            print("No checkpoints were restored!!")
        print("\nInitialized SNDCGAN successfully!\n")

    # -- checkpoints: CheckpointManager(max_to_keep=2) semantics with torch.save files ckpt-<epoch>.pt
    def _checkpoints(self):
        files = glob.glob(os.path.join(self.checkpoint_path, "ckpt-*.pt"))
        return sorted(files, key=lambda f: int(os.path.basename(f).split("-")[-1].split(".")[0]))

    def _latest_checkpoint(self):
        c = self._checkpoints()
        return c[-1] if c else None

    def _save(self, epoch):
        G, D = self.gen_model.arena, self.disc_model.arena
        state = {"epoch": epoch, "g_w": G.w.cpu(), "g_m": G.m.cpu(), "g_v": G.v.cpu(), "d_w": D.w.cpu(),
                 "d_m": D.m.cpu(), "d_v": D.v.cpu(), "g_iterations": self.g_optimizer.iterations,
                 "d_iterations": self.d_optimizer.iterations, "step_count": self.step_count}
        if self.disc_model.spectral_norm:
            state["sn_u"] = [u.cpu() for u in self.disc_model.sn_u]
        torch.save(state, os.path.join(self.checkpoint_path, f"ckpt-{epoch}.pt"))
        for old in self._checkpoints()[:-2]:
            os.remove(old)

    def _restore(self, path):
        st = torch.load(path, map_location="cpu")
        G, D = self.gen_model.arena, self.disc_model.arena
        for dst, key in ((G.w, "g_w"), (G.m, "g_m"), (G.v, "g_v"), (D.w, "d_w"), (D.m, "d_m"), (D.v, "d_v")):
            dst.copy_(st[key])
        self.g_optimizer.iterations, self.d_optimizer.iterations = st["g_iterations"], st["d_iterations"]
        self.step_count = st["step_count"]
        if self.disc_model.spectral_norm and "sn_u" in st:
            for u, s in zip(self.disc_model.sn_u, st["sn_u"]):
                u.copy_(s)
        self.gen_model._packed_dirty = True
        self.disc_model._packed_dirty = True

    def display_examples(self, samples, number_of_images, info_text):
        """:228-238 writes a matplotlib PDF; headless here: the preview strip goes to a PNG (Pillow)."""
        try:
            from PIL import Image
            strip = (samples.clamp(0, 1) * 255).to(torch.uint8).cpu().numpy()
            Image.fromarray(np.concatenate(list(strip[:number_of_images]), axis=1)).save(self.live_preview_file)
        except Exception as e:  # previews are a convenience, never fatal
            print(f"(preview not written: {e})")

    def train(self, num_epochs, checkpoint_frequency):
        """SNDCGAN.py:272-335 — same loop, logging keys and checkpoint cadence."""
        start_time = time()
        local_losses = {"epoch": [], "avg_g_loss": [], "avg_d_loss": [], "d_real": [], "d_fake": []}
        dev = self.device
        for epoch in range(self.start_epoch, num_epochs):
            epoch_losses, epoch_d_vals = [], []
            batch_start_time = time()
            for i, (input_real, _) in enumerate(self.train_ds):
                if input_real.shape[0] != self.batch_size:
                    continue   # kernels are shaped for the configured batch; the ragged last batch is skipped
                input_real = input_real.to(dev, non_blocking=True)
                input_z = torch.rand((self.batch_size, self.z_size), device=dev) * 2.0 - 1.0        # :283
                out = self.train_step(input_real, input_z)                                          # :284
                epoch_losses.append(torch.stack(out[:4]))
                epoch_d_vals.append(torch.stack((torch.sigmoid(out[4]).mean(), torch.sigmoid(out[5]).mean())))
                print(f"\r>Batch {i:03d}, passed time: {strftime('%M:%S', gmtime(time() - batch_start_time))}",
                      end="", flush=True)
            if not epoch_losses:
                raise RuntimeError("dataset yielded no full batch")
            avg_losses = torch.stack(epoch_losses).mean(dim=0).cpu().tolist()     # one host sync per epoch
            for key, val in zip(("epoch", "avg_g_loss", "avg_d_loss", "d_real", "d_fake"), [epoch] + avg_losses):
                local_losses[key].append(val)
            epoch_duration = strftime("%H:%M:%S", gmtime(time() - start_time))
            info_text = "Epoch {:04d} | ET {} min | Avg Losses G/D {:.4f}/{:.4f} [D-Real: {:.4f} D-Fake {:.4f}]".format(
                epoch, epoch_duration, *avg_losses)
            print("\r" + info_text)
            n_prev = 3
            fixed_z = torch.rand((n_prev, self.z_size), device=dev) * 2.0 - 1.0
            self.display_examples(create_samples(self.gen_model, fixed_z, n_prev), n_prev, info_text)
            if epoch % checkpoint_frequency == 0:                                                   # :317
                self._save(epoch)
                for key, val in local_losses.items():
                    self.losses[key] = self.losses[key] + val
                with open(self.losses_file, mode="wb") as f:
                    pickle.dump(self.losses, f)
                local_losses = {"epoch": [], "avg_g_loss": [], "avg_d_loss": [], "d_real": [], "d_fake": []}


def build_parser():
    """sndcgan/Trainer.py:12-31 flag for flag; extras after the blank line default to the reference's behaviour."""
    p = argparse.ArgumentParser(description="GAN Trainer to generate landscape images.")
    p.add_argument("bSize", type=int, help="Batch Size to use.")
    p.add_argument("epochs", type=int, help="Number of epochs to train.")
    p.add_argument("-cf", "--checkpointFrequency", type=int, dest="ckptFreq", default=5)
    p.add_argument("-d", "--directory", type=str, dest="dirPath", default="training")
    p.add_argument("-x", "--data", type=str, dest="data", default="dataset")
    p.add_argument("-r", "--dropout", type=float, dest="dropout", default=0.5)
    p.add_argument("-ld", "--learnRateDisc", type=float, dest="learnRateDisc", default=0.0002)
    p.add_argument("-lg", "--learnRateGen", type=float, dest="learnRateGen", default=0.0002)
    p.add_argument("-lo", "--liveOutput", type=str, dest="liveOutput", default="live")
    p.add_argument("-ct", "--continue", dest="continue_", action="store_true", default=False)

    p.add_argument("--spectral-norm", action="store_true", help="spectral normalisation of D (not in the reference)")
    p.add_argument("--loss", choices=["bce", "hinge"], default="bce")
    p.add_argument("--seed", type=int, default=62)
    return p


def main(argv=None):
    args = build_parser().parse_args(argv)
    sndcgan = SNDCGAN(args.dirPath, args.data, args.bSize, args.dropout, args.learnRateDisc, args.learnRateGen,
                      args.liveOutput, args.continue_, image_size, z_size, spectral_norm=args.spectral_norm,
                      loss=args.loss, seed=args.seed)
    sndcgan.train(args.epochs + 1, args.ckptFreq)


if __name__ == "__main__":
    main()
