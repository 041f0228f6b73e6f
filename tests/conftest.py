import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE_DIR = '/root/reference'


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a real B200 (run with -m gpu on the GPU box)')


def have_reference():
    return os.path.isdir(os.path.join(REFERENCE_DIR, 'AlexCap'))


@pytest.fixture(scope='session')
def reference_modules():
    """Import the UNMODIFIED reference modules (build container only).  One process-local shim:
    Decoder.forward hard-codes .cuda() on its output buffers (AttentionLanguageModule.py:70-71)."""
    if not have_reference():
        pytest.skip('/root/reference not present on this machine')
    import torch
    if not torch.cuda.is_available():
        torch.Tensor.cuda = lambda self, *a, **k: self
    sys.path.insert(0, REFERENCE_DIR)
    from AlexCap.AttentionLanguageModule import Decoder
    from AlexCap.LanguageModule import LanguageModel
    from AlexCap.CustomLoss import CustomCrossEntropyLoss
    return dict(Decoder=Decoder, LanguageModel=LanguageModel, CustomCrossEntropyLoss=CustomCrossEntropyLoss)


@pytest.fixture(params=[3, 1, 0], ids=['bf16x3', 'tcgen05', 'simt'])
def gemm_engine(request):
    """Run a GPU test once per GEMM engine: 3 = tcgen05 bf16x3 (the training default), 1 = tcgen05/TMEM/TMA 3xTF32, 0 = fp32 CUDA-core tiles."""
    from imagecaptioning_b200 import functional as F_
    F_.set_gemm_engine(request.param)
    yield request.param
    F_.set_gemm_engine(3)
