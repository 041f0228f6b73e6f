"""ORACLE helper (test infrastructure): the synthetic inputs of SURVEY.md section 8(d), shared by
tests, golden-vector generation and the bench so every arm sees identical tensors."""
import torch


def synth_inputs(B, L, D, V, T, seed=123, full_length=False, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    feat = torch.randn(B, L, D, generator=g).relu_().to(dtype)
    if full_length:
        lens = torch.full((B,), T, dtype=torch.long)
    else:
        lens = torch.randint(max(T // 2, 1), T + 1, (B,), generator=g)
    cap = torch.randint(1, V + 1, (B, T), generator=g, dtype=torch.int64)
    cap = cap * (torch.arange(T).unsqueeze(0) < lens.unsqueeze(1))
    return feat, cap.to(torch.int32)
