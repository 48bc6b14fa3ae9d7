"""Seeded synthetic models / batches shared by the golden generator and the parity tests.  TEST INFRASTRUCTURE ONLY.
Nothing here imports the reference."""

import torch


def gpt2_model_and_batches():
    """BASELINE.json configs[0]: random-init GPT-2 small from its config (no checkpoint), fp32, 4 x [4, 512] tokens."""
    from transformers import GPT2Config, GPT2LMHeadModel

    torch.manual_seed(0)
    model = GPT2LMHeadModel(GPT2Config()).eval()
    gen = torch.Generator().manual_seed(1234)
    batches = [torch.randint(0, 50257, (4, 512), generator=gen) for _ in range(4)]
    return model, batches
