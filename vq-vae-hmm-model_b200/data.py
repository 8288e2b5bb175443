"""Batch boundary and training call site, mirroring VQ_VAE_HMM_fixed.py:8-29, :145-179 and
src/utils/data.py:7-44 (same names, arguments and semantics).  Host logic only; the hot
path behind ``model.compute_loss`` is in libvhmm.so."""
from __future__ import annotations

import random

import torch
import torch.optim as optim

device = torch.device('cuda' if torch.cuda.is_available() else 'cpu')  # VQ_VAE_HMM_fixed.py:8


class RandomChunkDataset(torch.utils.data.Dataset):
    """VQ_VAE_HMM_fixed.py:10-29: a random chunk (length min_len..min(max_len, S)) of a random
    sequence per item; ``__len__`` is fixed at 1000; Python ``random`` drives the sampling."""

    def __init__(self, x_sequences, u_sequences, min_len=20, max_len=200):
        self.x_seqs = x_sequences
        self.u_seqs = u_sequences
        self.min_len = min_len
        self.max_len = max_len

    def __len__(self):
        return 1000

    def __getitem__(self, idx):
        i = random.randint(0, len(self.x_seqs) - 1)
        xs, us = self.x_seqs[i], self.u_seqs[i]
        S = xs.shape[1]
        n = random.randint(self.min_len, min(self.max_len, S))
        s = random.randint(0, S - n)
        return xs[:, s:s + n], us[:, s:s + n], n


def collate_fn(batch, device=None):
    """Zero-pad to the batch max length; lengths stay an int64 CPU tensor
    (VQ_VAE_HMM_fixed.py:164-179).  Accepts both reference arities: ``collate_fn(batch)`` uses the
    module-level ``device`` (root module), ``collate_fn(batch, device)`` is the src/utils/data.py:29
    form that training_pipeline/train.py:118 and evaluation/evaluate.py:72 call.
    Padding is assembled on the host and moved with one copy per tensor instead of 2B slice
    copies into device memory."""
    dev = globals()["device"] if device is None else torch.device(device)
    lengths = torch.tensor([item[2] for item in batch], dtype=torch.long)
    T = int(lengths.max().item())
    B = len(batch)
    C = batch[0][0].shape[0]
    U = batch[0][1].shape[0]
    pin = dev.type == "cuda"
    xh = torch.zeros(B, C, T, pin_memory=pin)
    uh = torch.zeros(B, U, T, pin_memory=pin)
    for i, (x, u, L) in enumerate(batch):
        xh[i, :, :L] = x
        uh[i, :, :L] = u
    return xh.to(dev, non_blocking=True), uh.to(dev, non_blocking=True), lengths


def train_model(model, dataloader, num_epochs=10, lr=1e-3):
    """VQ_VAE_HMM_fixed.py:145-162: Adam, beta warm-up min(1, 2(ep+1)/E), mean loss printed."""
    opt = optim.Adam(model.parameters(), lr=lr)
    model.train()
    for ep in range(num_epochs):
        epoch_loss = 0
        beta = min(1.0, 2.0 * (ep + 1) / num_epochs)
        for x, u, lengths in dataloader:
            opt.zero_grad()
            loss = model.compute_loss(x, u, lengths, beta)
            loss.backward()
            opt.step()
            epoch_loss += loss.item()
        print(f"Epoch {ep+1}/{num_epochs}, Loss: {epoch_loss/len(dataloader):.4f}")
    return model


def save_checkpoint(model, optimizer, epoch, loss, path):
    """src/utils/data.py:47-53."""
    torch.save({'epoch': epoch, 'model_state_dict': model.state_dict(),
                'optimizer_state_dict': optimizer.state_dict(), 'loss': loss}, path)


def load_checkpoint(model, optimizer, path):
    """src/utils/data.py:56-60."""
    ck = torch.load(path)
    model.load_state_dict(ck['model_state_dict'])
    optimizer.load_state_dict(ck['optimizer_state_dict'])
    return ck['epoch'], ck['loss']
