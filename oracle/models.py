"""Seeded synthetic models / batches shared by the golden generator and the parity tests.  TEST INFRASTRUCTURE ONLY.
Nothing here imports the reference.

Values come from an integer hash (exact int64 tensor arithmetic), not from torch's CPU normal_/randn: those sample
through vectorised math whose results differ between CPU models, and the golden vectors are made in one container and
checked on another machine."""

import torch


def hash_uniform(n: int, seed: int) -> torch.Tensor:
    """n float64 values in [0, 1), identical on every machine."""
    x = (torch.arange(n, dtype=torch.int64) + seed * 0x9E3779B1) & 0xFFFFFFFF
    for _ in range(2):
        x = (((x >> 16) ^ x) * 0x45D9F3B) & 0xFFFFFFFF
    x = (x >> 16) ^ x
    return x.double() / 4294967296.0


@torch.no_grad()
def deterministic_fill_(model, seed: int = 0):
    for k, (name, p) in enumerate(model.named_parameters()):
        u = hash_uniform(p.numel(), seed * 1000 + k).reshape(p.shape)
        if p.dim() >= 2:
            v = (u - 0.5) * (2 * 0.02 * 1.7320508)      # std 0.02
        elif name.endswith("weight"):
            v = 1.0 + (u - 0.5) * 0.2                    # norm gains
        else:
            v = (u - 0.5) * 0.02                         # biases
        p.copy_(v.to(p.dtype))
    return model


def fingerprint(model) -> torch.Tensor:
    """Order-independent exact checksum per parameter: the int64 sum of the fp32 bit patterns."""
    return torch.stack([p.detach().float().contiguous().view(torch.int32).long().sum() for p in model.parameters()])


def gpt2_model_and_batches():
    """BASELINE.json configs[0]: random-init GPT-2 small from its config (no checkpoint), fp32, 4 x [4, 512] tokens."""
    from transformers import GPT2Config, GPT2LMHeadModel

    model = deterministic_fill_(GPT2LMHeadModel(GPT2Config()).eval(), seed=0)
    batches = [(hash_uniform(4 * 512, 77 + i) * 50257).long().reshape(4, 512) for i in range(4)]
    return model, batches
