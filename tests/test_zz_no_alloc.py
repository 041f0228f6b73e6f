"""-m gpu, collected LAST: alexcap_cross_entropy_static (include/alexcap.h, csrc/no_alloc.cu) -- the cross-entropy entry point
with its scratch in a file-scope __device__ array instead of a caller buffer -- through the C ABI against the existing
alexcap_cross_entropy on the same inputs, on two calls in a row (the second call meets the first call's leftovers in the
shared array, with a different N so the n_valid slot moves), plus the size limit.  No counterpart in the reference."""
import pytest
import torch

from tests.gpu_util import DEV
from tests.util import rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-5          # the tolerance of tests/test_gpu_kernels.py::test_cross_entropy_value_and_grad


def _case(N, C, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(N, C, generator=g) * 3
    t = torch.randint(0, C, (N,), generator=g)
    t[::4] = 0                                            # ignored rows
    return x.to(DEV), t.to(torch.int32).to(DEV)


def test_static_scratch_cross_entropy_equals_existing_entry_point_twice_in_a_row():
    from imagecaptioning_b200 import functional as F_
    lib = F_.ensure_device(DEV)
    for N, C, seed in [(300, 1003, 1), (37, 53, 2)]:
        xv, t = _case(N, C, seed)
        # existing entry point (caller-provided scratch)
        out_ref, dl_ref = F_.cross_entropy(xv, t)
        # new entry point: every buffer passed in, nothing allocated inside the call
        out = torch.full((2,), float('nan'), device=DEV)
        dl = torch.full_like(xv, float('nan'))
        before = torch.cuda.memory_allocated()
        F_.cross_entropy_static(xv, t, out, dl)
        assert torch.cuda.memory_allocated() == before
        torch.cuda.synchronize()
        assert abs(float(out[0]) - float(out_ref[0])) <= TOL * max(1.0, abs(float(out_ref[0])))
        assert float(out[1]) == float(out_ref[1]) == float((t != 0).sum())
        assert rel_err(dl, dl_ref) < TOL
        # and against torch's definition of CustomLoss.py:7 (the oracle of the existing test)
        xr = xv.double().cpu().requires_grad_(True)
        ref = torch.nn.functional.cross_entropy(xr, t.long().cpu(), ignore_index=0, label_smoothing=0.1)
        ref.backward()
        assert abs(float(out[0]) - float(ref.detach())) < TOL * max(1.0, abs(float(ref.detach())))
        assert rel_err(dl, xr.grad) < TOL
    # value only (dlogits NULL) after the gradient calls
    out2 = torch.empty(2, device=DEV)
    F_.cross_entropy_static(xv, t, out2, None)
    assert abs(float(out2[0]) - float(out_ref[0])) <= TOL * max(1.0, abs(float(out_ref[0])))
    assert lib.alexcap_cross_entropy_static_max_rows() == 65536


def test_static_scratch_cross_entropy_rejects_sizes_beyond_the_array():
    from imagecaptioning_b200 import functional as F_
    from imagecaptioning_b200._lib import AlexcapError
    lib = F_.ensure_device(DEV)
    N = lib.alexcap_cross_entropy_static_max_rows() + 1
    x = torch.zeros(N, 4, device=DEV)
    t = torch.ones(N, dtype=torch.int32, device=DEV)
    out = torch.empty(2, device=DEV)
    with pytest.raises(AlexcapError, match='exceeds the static scratch'):
        F_.cross_entropy_static(x, t, out, None)
    # the limit itself is accepted
    F_.cross_entropy_static(x[:N - 1], t[:N - 1], out, None)
    torch.cuda.synchronize()
    assert float(out[1]) == N - 1
    assert abs(float(out[0]) - torch.log(torch.tensor(4.0)).item()) < 1e-5
