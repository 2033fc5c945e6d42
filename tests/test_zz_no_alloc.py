"""nbody_gpu_forces_static -- the entry point for callers that may not allocate device memory -- against the
handle-based nbody_gpu_forces, through the C ABI.

Some users of the library cannot call cudaMalloc: a plug-in inside an application whose own pool claimed the GPU's
memory at start-up, or a loop that cannot take the device-wide stall a cudaMalloc/cudaFree may cause.  For them the
library keeps the directed kernel's per-chunk partial sums in a fixed 16 MiB array of its own, and the caller passes
device arrays it already has.  Here torch plays that caller: it owns the input and output buffers, nothing else.

The file is named test_zz_* so that pytest collects it last: a failure here must not hide other tests behind -x.
It passes alone in a fresh process (the first call does the per-device address look-up) and after the other tests.
"""
import numpy as np
import pytest


def _interleave(a, b):
    out = np.empty(2 * len(a))
    out[0::2] = a
    out[1::2] = b
    return out


def _reference_forces(nb, b):
    """The existing entry point: a handle in fast mode, directed evaluation, creation order, default chunks."""
    g = nb.NBodyGPU(b.x, b.y, b.xv, b.yv, b.mass, b.radius, mode=nb.MODE_FAST, device=0, symmetric=-1, sort=-1)
    try:
        assert not g.describe()["symmetric"]
        return g.forces()
    finally:
        g.close()


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1000, 20011, 32768])
def test_static_scratch_forces_match_the_handle_twice_in_a_row(nb, oracle, n):
    import torch
    torch.cuda.set_device(0)
    b = oracle.init_bodies(n)
    want_xf, want_yf = _reference_forces(nb, b)
    fmax = max(np.abs(want_xf).max(), np.abs(want_yf).max())

    pos = torch.from_numpy(_interleave(b.x, b.y)).cuda()
    mr = torch.from_numpy(_interleave(b.mass, b.radius)).cuda()
    out = torch.full((2 * n,), float("nan"), dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream

    # second round on moved bodies: the shared scratch still holds the first call's partial sums
    x2 = np.ascontiguousarray(b.x[::-1]) * 0.5 + 3.25
    y2 = np.ascontiguousarray(b.y[::-1]) * 0.75 + 1.5
    b2 = b.copy()
    b2.x[:], b2.y[:] = x2, y2
    want2_xf, want2_yf = _reference_forces(nb, b2)
    fmax2 = max(np.abs(want2_xf).max(), np.abs(want2_yf).max())

    before = torch.cuda.memory_allocated()
    nb.forces_static(n, pos.data_ptr(), mr.data_ptr(), out.data_ptr(), stream)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    # fast-mode force tolerance of this repository's tests: 1e-8 * max|F| (tests/test_gpu_parity.py); the two
    # entry points run the same kernels over the same chunks, so they agree to the bit as well
    assert np.abs(got[0::2] - want_xf).max() <= 1e-8 * fmax
    assert np.abs(got[1::2] - want_yf).max() <= 1e-8 * fmax
    assert np.array_equal(got[0::2].view(np.uint64), want_xf.view(np.uint64))
    assert np.array_equal(got[1::2].view(np.uint64), want_yf.view(np.uint64))

    pos.copy_(torch.from_numpy(_interleave(x2, y2)))
    out.fill_(float("nan"))
    nb.forces_static(n, pos.data_ptr(), mr.data_ptr(), out.data_ptr(), stream)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    assert np.abs(got[0::2] - want2_xf).max() <= 1e-8 * fmax2
    assert np.abs(got[1::2] - want2_yf).max() <= 1e-8 * fmax2
    assert np.array_equal(got[0::2].view(np.uint64), want2_xf.view(np.uint64))
    assert np.array_equal(got[1::2].view(np.uint64), want2_yf.view(np.uint64))
    assert torch.cuda.memory_allocated() == before   # the caller's allocator saw no request either


@pytest.mark.gpu
def test_static_scratch_rejects_what_it_cannot_hold(nb):
    import torch
    torch.cuda.set_device(0)
    buf = torch.zeros(64, dtype=torch.float64, device="cuda")
    host = np.zeros(64)
    with pytest.raises(nb.NBodyGPUError, match="NBODY_GPU_STATIC_MAX_BODIES"):
        nb.forces_static(nb.STATIC_MAX_BODIES + 1, buf.data_ptr(), buf.data_ptr(), buf.data_ptr())
    with pytest.raises(nb.NBodyGPUError, match="at least one body"):
        nb.forces_static(0, buf.data_ptr(), buf.data_ptr(), buf.data_ptr())
    with pytest.raises(nb.NBodyGPUError, match="NULL"):
        nb.forces_static(16, buf.data_ptr(), None, buf.data_ptr())
    with pytest.raises(nb.NBodyGPUError, match="16-byte"):
        nb.forces_static(16, buf.data_ptr() + 8, buf.data_ptr(), buf.data_ptr())
    with pytest.raises(nb.NBodyGPUError, match="not device memory"):
        nb.forces_static(16, host.ctypes.data // 16 * 16, buf.data_ptr(), buf.data_ptr())


def test_static_scratch_limit_is_the_header_s(nb):
    """CPU: the binding's constants are the header's, and the limit is refused before any CUDA call."""
    import os
    import re
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "nbody_gpu.h")).read()
    assert int(re.search(r"#define NBODY_GPU_STATIC_MAX_BODIES (\d+)", hdr).group(1)) == nb.STATIC_MAX_BODIES
    assert re.search(r"#define NBODY_GPU_STATIC_SCRATCH_BYTES \(16u << 20\)", hdr) and nb.STATIC_SCRATCH_BYTES == 16 << 20
    with pytest.raises(nb.NBodyGPUError, match="exceeds NBODY_GPU_STATIC_MAX_BODIES"):
        nb.forces_static(nb.STATIC_MAX_BODIES + 1, 16, 16, 16)
