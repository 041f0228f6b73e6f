"""-m gpu: the tcgen05 / TMEM / TMA GEMM engine (3xTF32) against fp64 math, every operand-major combination,
edge shapes, split-K, epilogues -- and the full training step running on it."""
import pytest
import torch

from tests.util import rel_err

pytestmark = pytest.mark.gpu
DEV = 'cuda'


@pytest.fixture(scope='module')
def F_():
    from imagecaptioning_b200 import functional
    functional.ensure_device(DEV)
    return functional


def _mk(ta, tb, M, N, K, g, lda_pad=0, ldb_pad=0):
    A = torch.randn((K, M + lda_pad) if ta else (M, K + lda_pad), generator=g)
    B = torch.randn((N, K + ldb_pad) if tb else (K, N + ldb_pad), generator=g)
    Av = A[:, :M] if ta else A[:, :K]
    Bv = B[:, :K] if tb else B[:, :N]
    ref = (Av.double().T if ta else Av.double()) @ (Bv.double().T if tb else Bv.double())
    return A, B, ref


@pytest.mark.parametrize('ta,tb', [(False, True), (False, False), (True, False), (True, True)])
@pytest.mark.parametrize('M,N,K', [(128, 128, 32), (128, 64, 64), (256, 256, 256), (37, 53, 29), (130, 200, 100),
                                   (128, 2048, 1024), (2176, 515, 96), (1, 7, 3)])
def test_gemm_tc_layouts_and_edges(F_, ta, tb, M, N, K):
    g = torch.Generator().manual_seed(M * 1000 + N + K)
    pad_a = (-M) % 4 if ta else (-K) % 4          # TMA needs ld % 4 == 0
    pad_b = (-K) % 4 if tb else (-N) % 4
    A, B, ref = _mk(ta, tb, M, N, K, g, pad_a, pad_b)
    out = F_.gemm(A.to(DEV), B.to(DEV), transa=ta, transb=tb, M=M, N=N, K=K, lda=A.shape[1], ldb=B.shape[1], engine=1)
    torch.cuda.synchronize()
    assert rel_err(out, ref) < 2e-5, rel_err(out, ref)


@pytest.mark.parametrize('splits', [2, 3, 16])
def test_gemm_tc_split_k(F_, splits):
    g = torch.Generator().manual_seed(splits)
    M, N, K = 128, 512, 2048
    A, B, ref = _mk(False, False, M, N, K, g)
    bias = torch.randn(N, generator=g)
    C0 = torch.randn(M, N, generator=g)
    out = F_.gemm(A.to(DEV), B.to(DEV), transa=False, transb=False, bias=bias.to(DEV), engine=splits)
    assert rel_err(out, ref + bias.double()) < 2e-5
    Cd = C0.to(DEV)
    F_.gemm(A.to(DEV), B.to(DEV), transa=False, transb=False, out=Cd, accumulate=True, alpha=0.5, engine=splits)
    assert rel_err(Cd, 0.5 * ref + C0.double()) < 2e-5


def test_gemm_tc_epilogues(F_):
    g = torch.Generator().manual_seed(5)
    M, N, K = 70, 45, 36
    A, B, ref = _mk(False, True, M, N, K, g)
    b1, b2 = torch.randn(N, generator=g), torch.randn(N, generator=g)
    C0 = torch.randn(M, N + 3, generator=g)
    Cd = C0.to(DEV)
    F_.gemm(A.to(DEV), B.to(DEV), out=Cd, ldc=N + 3, M=M, N=N, K=K, bias=b1.to(DEV), bias2=b2.to(DEV), accumulate=True,
            act=1, alpha=0.5, engine=1)
    want = torch.tanh(0.5 * ref + b1.double() + b2.double() + C0[:, :N].double())
    assert rel_err(Cd[:, :N], want) < 2e-5
    assert torch.equal(Cd[:, N:].cpu(), C0[:, N:])
    out = F_.gemm(A.to(DEV), B.to(DEV), act=2, engine=1)
    assert rel_err(out, ref.clamp_min(0)) < 2e-5


def test_gemm_tc_accuracy_is_fp32_grade(F_):
    """3xTF32 must beat single-pass TF32 by orders of magnitude: K=4096 dot products of O(1) numbers."""
    g = torch.Generator().manual_seed(9)
    A, B, ref = _mk(False, True, 256, 256, 4096, g)
    out = F_.gemm(A.to(DEV), B.to(DEV), engine=1)
    simt = F_.gemm(A.to(DEV), B.to(DEV), engine=0)
    e_tc, e_simt = rel_err(out, ref), rel_err(simt, ref)
    assert e_tc < 5e-6, (e_tc, e_simt)


@pytest.mark.parametrize('ta,tb', [(False, True), (False, False), (True, False), (True, True)])
@pytest.mark.parametrize('M,N,K', [(128, 128, 32), (128, 64, 64), (256, 256, 256), (37, 53, 29), (130, 200, 100),
                                   (128, 2048, 1024), (2176, 515, 96)])
def test_gemm_bf16_engine_layouts(F_, ta, tb, M, N, K):
    """bf16-operand engine (dims.dtype = ALEXCAP_BF16 path): exactly the fp64 product of the bf16-ROUNDED operands up to
    fp32 accumulation error, for every operand-major combination (this pins the in-kernel convert / re-layout)."""
    g = torch.Generator().manual_seed(M * 1000 + N + K + 7)
    pad_a = (-M) % 4 if ta else (-K) % 4
    pad_b = (-K) % 4 if tb else (-N) % 4
    A, B, _ = _mk(ta, tb, M, N, K, g, pad_a, pad_b)
    Av = (A[:, :M] if ta else A[:, :K]).bfloat16().double()
    Bv = (B[:, :K] if tb else B[:, :N]).bfloat16().double()
    ref = (Av.T if ta else Av) @ (Bv.T if tb else Bv)
    bias = torch.randn(N, generator=g)
    out = F_.gemm(A.to(DEV), B.to(DEV), transa=ta, transb=tb, M=M, N=N, K=K, lda=A.shape[1], ldb=B.shape[1],
                  bias=bias.to(DEV), engine=100)
    assert rel_err(out, ref + bias.double()) < 2e-5, rel_err(out, ref + bias.double())
    out = F_.gemm(A.to(DEV), B.to(DEV), transa=ta, transb=tb, M=M, N=N, K=K, lda=A.shape[1], ldb=B.shape[1], engine=103)
    assert rel_err(out, ref) < 2e-5            # split-K forced to 3


# ---------------------------------------------------------------------------------------------- pre-split engine (gemm_ps.cu)
PS_SHAPES = [(128, 128, 64), (256, 256, 256), (37, 53, 29), (130, 200, 100), (384, 1000, 520), (2176, 515, 96),
             (2300, 1100, 200), (1, 7, 3)]


@pytest.mark.parametrize('ta,tb', [(False, True), (False, False), (True, False), (True, True)])
@pytest.mark.parametrize('M,N,K', PS_SHAPES)
def test_gemm_presplit_bf16_layouts(F_, ta, tb, M, N, K):
    """Persistent TMA -> tcgen05 kernel over operands split in HBM, bf16 mode: the fp64 product of the bf16-ROUNDED operands
    up to fp32 accumulation error, for every operand-major combination (pins the TMA boxes and UMMA descriptors; 2300x1100
    gives 162 tiles > 148 CTAs, so some CTAs run two work items through both TMEM accumulators)."""
    g = torch.Generator().manual_seed(M * 1000 + N + K + 11)
    pad_a = (-M) % 4 if ta else (-K) % 4
    pad_b = (-K) % 4 if tb else (-N) % 4
    A, B, _ = _mk(ta, tb, M, N, K, g, pad_a, pad_b)
    Av = (A[:, :M] if ta else A[:, :K]).bfloat16().double()
    Bv = (B[:, :K] if tb else B[:, :N]).bfloat16().double()
    ref = (Av.T if ta else Av) @ (Bv.T if tb else Bv)
    bias = torch.randn(N, generator=g)
    out = F_.gemm(A.to(DEV), B.to(DEV), transa=ta, transb=tb, M=M, N=N, K=K, lda=A.shape[1], ldb=B.shape[1],
                  bias=bias.to(DEV), engine=400)
    assert rel_err(out, ref + bias.double()) < 2e-5, rel_err(out, ref + bias.double())
    if K >= 128:
        out = F_.gemm(A.to(DEV), B.to(DEV), transa=ta, transb=tb, M=M, N=N, K=K, lda=A.shape[1], ldb=B.shape[1], engine=402)
        assert rel_err(out, ref) < 2e-5            # split-K forced to 2


@pytest.mark.parametrize('ta,tb', [(False, True), (False, False), (True, False), (True, True)])
@pytest.mark.parametrize('M,N,K', PS_SHAPES)
def test_gemm_presplit_bf16x3_is_fp32_grade(F_, ta, tb, M, N, K):
    g = torch.Generator().manual_seed(M * 1000 + N + K + 13)
    pad_a = (-M) % 4 if ta else (-K) % 4
    pad_b = (-K) % 4 if tb else (-N) % 4
    A, B, ref = _mk(ta, tb, M, N, K, g, pad_a, pad_b)
    C0 = torch.randn(M, N, generator=g)
    Cd = C0.to(DEV)
    F_.gemm(A.to(DEV), B.to(DEV), transa=ta, transb=tb, M=M, N=N, K=K, lda=A.shape[1], ldb=B.shape[1], out=Cd,
            accumulate=True, alpha=0.5, engine=300)
    assert rel_err(Cd, 0.5 * ref + C0.double()) < 2e-5
    if K >= 192:
        out = F_.gemm(A.to(DEV), B.to(DEV), transa=ta, transb=tb, M=M, N=N, K=K, lda=A.shape[1], ldb=B.shape[1], engine=303)
        assert rel_err(out, ref) < 2e-5


def test_gemm_presplit_epilogues_and_long_k(F_):
    g = torch.Generator().manual_seed(21)
    M, N, K = 300, 200, 4100
    A, B, ref = _mk(False, True, M, N, K, g)
    b1, b2 = torch.randn(N, generator=g), torch.randn(N, generator=g)
    out = F_.gemm(A.to(DEV), B.to(DEV), bias=b1.to(DEV), bias2=b2.to(DEV), act=1, alpha=0.01, engine=300)
    assert rel_err(out, torch.tanh(0.01 * ref + b1.double() + b2.double())) < 1e-4   # |z| up to ~5, output scale 1
    out = F_.gemm(A.to(DEV), B.to(DEV), act=2, engine=300)
    assert rel_err(out, ref.clamp_min(0)) < 2e-5
    out = F_.gemm(A.to(DEV), B.to(DEV), engine=300)
    e3, e1 = rel_err(out, ref), rel_err(F_.gemm(A.to(DEV), B.to(DEV), engine=400), ref)
    assert e3 < 1e-5 and e1 > 100 * e3, (e3, e1)     # hi + lo keeps ~17 significand bits per operand


@pytest.mark.parametrize('engine', [0, 1, 200, 300])
@pytest.mark.parametrize('M,N,K', [(300, 260, 128), (2176, 515, 96)])
def test_gemm_exp2c_activation_all_engines(F_, engine, M, N, K):
    """act = 3 (ACT_EXP2C, common.cuh): C = 2^(clamp(2 log2(e) (A B^T + bias), +-60)) = e^(2x) -- the form the hoisted
    W(features) GEMM stores for the streaming attention kernels -- on the CUDA-core, 3xTF32, bf16x3 and pre-split engines,
    aligned (vector epilogue) and odd N (scalar epilogue), including the clamp."""
    g = torch.Generator().manual_seed(M + N + engine)
    A = torch.randn(M, K, generator=g) / K ** 0.5 * 3
    B = torch.randn(N, K, generator=g)
    bias = torch.randn(N, generator=g)
    A[0] *= 40                                                       # a row whose pre-activations leave the clamp range
    z = 2.8853900817779268 * (A.double() @ B.double().T + bias.double())
    ref = torch.exp2(z.clamp(-60, 60))
    out = F_.gemm(A.to(DEV), B.to(DEV), transa=False, transb=True, bias=bias.to(DEV), act=3, engine=engine).double().cpu()
    torch.cuda.synchronize()
    # rows 1..: element-wise relative (e^(2x) spans many decades; d e^(2x) / e^(2x) = 2 dx, the error of the pre-activation doubled)
    rel = ((out[1:] - ref[1:]).abs() / ref[1:]).max()
    assert float(rel) < (2e-3 if engine else 2e-4), float(rel)
    # row 0: far outside the clamp range -> exactly 2^+-60 wherever the pre-activation is clearly beyond it
    far = z[0].abs() > 70
    assert int(far.sum()) > N // 2
    assert torch.equal(out[0][far], ref[0][far])
