"""GPU test (-m gpu) of b200amg_dot_noalloc, the allocation-free entry point (include/b200amg.h): the dot product
with its partial sums in a library-owned static device array instead of a caller-provided work buffer.  It has no
counterpart in the reference; it is checked, through the C-ABI, against the existing entry point b200amg_dot.
Collected last on purpose (file name)."""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

RTOL = 1e-12            # the tolerance of the other norm / dot tests (the two entry points share kernels: expect bits)
MAX_N = 1 << 28


@pytest.fixture(scope='module')
def amg():
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    import pyamg_b200
    return pyamg_b200


def test_dot_noalloc_matches_dot_on_two_calls_in_a_row(amg):
    lib = amg._lib
    rng = np.random.RandomState(11)
    work = torch.empty(1024, dtype=torch.float64, device='cuda')
    outs = torch.zeros(4, dtype=torch.float64, device='cuda')
    stream = torch.cuda.current_stream().cuda_stream
    cases = []
    for n in (1 << 20, 1000003):              # two calls in a row share the static array: different n, different data
        x = torch.from_numpy(rng.rand(n) - 0.3).cuda()
        y = torch.from_numpy(rng.rand(n) - 0.6).cuda()
        cases.append((n, x, y))
    torch.cuda.synchronize()
    # every buffer exists before the calls: the entry point itself must not allocate
    before = torch.cuda.memory_allocated()
    for k, (n, x, y) in enumerate(cases):
        lib.call('b200amg_dot_noalloc', n, x.data_ptr(), y.data_ptr(), outs[2 * k:].data_ptr(), stream)
    for k, (n, x, y) in enumerate(cases):
        lib.call('b200amg_dot', n, x.data_ptr(), y.data_ptr(), outs[2 * k + 1:].data_ptr(), work.data_ptr(), stream)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    got = outs.cpu().numpy()
    for k, (n, x, y) in enumerate(cases):
        new, old = got[2 * k], got[2 * k + 1]
        host = float(np.dot(x.cpu().numpy(), y.cpu().numpy()))
        assert abs(new - old) <= RTOL * abs(old), (n, new, old)
        assert abs(new - host) <= 1e-11 * max(abs(host), 1.0), (n, new, host)


def test_dot_noalloc_small_and_empty(amg):
    lib = amg._lib
    work = torch.empty(1024, dtype=torch.float64, device='cuda')
    out = torch.full((2,), 7.0, dtype=torch.float64, device='cuda')
    x = torch.arange(1, 8, dtype=torch.float64, device='cuda')
    for n in (7, 1, 0):
        lib.call('b200amg_dot_noalloc', n, x.data_ptr(), x.data_ptr(), out.data_ptr(), 0)
        lib.call('b200amg_dot', n, x.data_ptr(), x.data_ptr(), out[1:].data_ptr(), work.data_ptr(), 0)
        torch.cuda.synchronize()
        a, b = out.cpu().numpy()
        assert a == b == float(sum(i * i for i in range(1, n + 1)))


def test_dot_noalloc_rejects_sizes_beyond_the_maximum(amg):
    lib = amg._lib
    x = torch.ones(8, dtype=torch.float64, device='cuda')
    out = torch.zeros(1, dtype=torch.float64, device='cuda')
    with pytest.raises(lib.B200AmgError, match='exceeds the maximum'):
        lib.call('b200amg_dot_noalloc', MAX_N + 1, x.data_ptr(), x.data_ptr(), out.data_ptr(), 0)
    raw = lib.load().b200amg_dot_noalloc(ctypes.c_longlong(MAX_N + 1), x.data_ptr(), x.data_ptr(), out.data_ptr(), 0)
    assert raw == -1                           # B200AMG_EINVAL, nothing launched
    torch.cuda.synchronize()
    assert float(out.item()) == 0.0
