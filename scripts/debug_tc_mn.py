"""Probe how tcgen05 reads MN-major shared-memory tiles (structured inputs reveal the element mapping).
Round-1 finding on B200: tf32 MN-major needs UMMA layout SWIZZLE_128B_BASE32B + TMA swizzle 128B_ATOM_32B,
LBO=4096, SBO=512 (all 4096/4096 probe entries correct).  Kept as a regression probe for the shipped constants."""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from imagecaptioning_b200 import functional as F_, _lib
lib = _lib.load()
dev = 'cuda'
M, N, K = 128, 64, 32

def probe_a():
    A = (torch.arange(K).view(K, 1) * 128 + torch.arange(M).view(1, M)).float()      # stored [K,M]
    B = torch.zeros(N, K); B[torch.arange(32), torch.arange(32)] = 1.0                 # [N,K] K-major
    C = F_.gemm(A.to(dev), B.to(dev), transa=True, transb=True, engine=1).cpu()
    v = C[:, :32].round().long()
    kk, mm = v // 128, v % 128
    ok = int(((kk == torch.arange(32).view(1, 32)) & (mm == torch.arange(M).view(M, 1))).sum())
    return ok, mm[:, 0].tolist(), kk[0].tolist(), mm[0].tolist(), kk[:, 0].tolist()[:40]

def probe_b():
    A = torch.zeros(M, K); A[torch.arange(32), torch.arange(32)] = 1.0                 # [M,K] K-major
    B = (torch.arange(K).view(K, 1) * 64 + torch.arange(N).view(1, N)).float()        # stored [K,N]
    C = F_.gemm(A.to(dev), B.to(dev), transa=False, transb=False, engine=1).cpu()
    v = C[:32].round().long()
    kk, nn = v // 64, v % 64
    ok = int(((kk == torch.arange(32).view(32, 1)) & (nn == torch.arange(N).view(1, N))).sum())
    return ok, nn[0].tolist(), kk[:, 0].tolist(), kk[0].tolist()[:40]

ok, *_ = probe_a(); print('A MN-major probe correct', ok, 'of', M * 32)
okb, *_ = probe_b(); print('B MN-major probe correct', okb, 'of', 32 * N)
