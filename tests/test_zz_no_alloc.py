"""GPU: macbf_loss_forward_noalloc, the loss reduction for callers that cannot allocate.

Some users of this library have no way to hand it device scratch: a host framework whose
allocator is frozen while it records or replays a step, a hard-real-time controller that must
stay out of the driver's allocator, or a plain C / Fortran simulator that owns no device
allocator at all.  For them the library offers one entry point whose partial-sum buffer is a
fixed array inside the .so.  This file checks that it returns exactly what the ordinary entry
point (caller-provided scratch) returns - the two run the same kernels, so the comparison is
bit-exact, tighter than the 1e-5 the other loss tests use - on two calls in a row with different
inputs (the array keeps the first call's partial sums; the second call must not see them), and
that a problem beyond the array's stated capacity is refused with a readable message.

Collected last on purpose (file name): a failure here cannot hide other tests under `-x`.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _problem(seed, B, N, K=12):
    from oracle import macbf_oracle as O
    s, g = O.synthetic_swarm(B, N, seed)
    gen = torch.Generator().manual_seed(seed)
    h = torch.randn(B, N, K, generator=gen) * 0.1
    hn = h + torch.randn(B, N, K, generator=gen) * 0.01
    u = torch.randn(B, N, 2, generator=gen)
    return [t.cuda().contiguous() for t in (s, g, h, hn, u)]


def _call_noalloc(L, s, g, idx, h, hn, u):
    B, N, K = h.shape
    mask = torch.full((B, N, K), 7, dtype=torch.uint8, device="cuda")
    scal = torch.full((B, 12), float("nan"), dtype=torch.float32, device="cuda")
    rc = L.macbf_loss_forward_noalloc(s.data_ptr(), g.data_ptr(), None, idx.data_ptr(), h.data_ptr(), hn.data_ptr(),
                                      u.data_ptr(), B, N, 0, K, 0.07, 2.0, 0.1, 1.0, mask.data_ptr(), scal.data_ptr(),
                                      torch.cuda.current_stream().cuda_stream)
    return rc, scal, mask


def test_noalloc_matches_scratch_entry_twice_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from macbf_b200 import lib, ops
    L = lib.load()
    # two different problems back to back: the second is smaller, so stale partial sums of the
    # first would still sit in the array beyond (and inside) the range the second one uses
    for seed, B, N in ((11, 6, 700), (12, 3, 130)):
        s, g, h, hn, u = _problem(seed, B, N)
        idx = ops.neighbor_topk(s, 12)
        ref_scal, ref_mask = ops.loss_forward(s, idx, h, hn, u, g)          # existing entry, torch-allocated scratch
        rc, scal, mask = _call_noalloc(L, s, g, idx, h, hn, u)
        assert rc == 0, L.macbf_last_error()
        torch.cuda.synchronize()
        assert torch.equal(mask, ref_mask)
        assert np.array_equal(scal.cpu().numpy(), ref_scal.cpu().numpy())
        assert np.isfinite(scal.cpu().numpy()).all()


def test_noalloc_rejects_problems_beyond_the_fixed_array():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from macbf_b200 import lib
    L = lib.load()
    # 4096 envs x 4096 agents x 12 slots -> 4096 * 48 chunks * 10 sums = 1,966,080 floats > 65,536
    assert L.macbf_loss_scratch_bytes(4096, 4096, 12) > 4 * 65536
    dummy = torch.zeros(16, device="cuda")
    p = dummy.data_ptr()
    rc = L.macbf_loss_forward_noalloc(p, p, None, p, p, p, p, 4096, 4096, 0, 12, 0.07, 2.0, 0.1, 1.0, p, p, None)
    assert rc != 0
    msg = L.macbf_last_error().decode()
    assert "too large" in msg and "macbf_loss_forward" in msg


def test_noalloc_rejects_misaligned_buffers_with_an_error_code():
    """A caller that carves its buffers out of one block can hand over a misaligned pointer; the kernels read states
    as float4 and goals / u as float2, so this has to come back as an error code, not as a GPU fault."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from macbf_b200 import lib, ops
    L = lib.load()
    s, g, h, hn, u = _problem(13, 2, 64)
    idx = ops.neighbor_topk(s, 12)
    B, N, K = h.shape
    mask = torch.zeros((B, N, K), dtype=torch.uint8, device="cuda")
    scal = torch.zeros((B, 12), dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for which in ("states", "goals", "u"):
        ptr = {"states": s.data_ptr(), "goals": g.data_ptr(), "u": u.data_ptr()}
        ptr[which] += 4
        for entry, extra in ((L.macbf_loss_forward_noalloc, ()),
                             (L.macbf_loss_forward, (scal.data_ptr(), 1 << 20))):   # scratch unused: rejected before
            args = [ptr["states"], ptr["goals"], None, idx.data_ptr(), h.data_ptr(), hn.data_ptr(), ptr["u"], B, N, 0, K,
                    0.07, 2.0, 0.1, 1.0, mask.data_ptr(), scal.data_ptr(), *extra, st]
            assert entry(*args) != 0
            assert "aligned" in L.macbf_last_error().decode()
    torch.cuda.synchronize()     # no sticky fault: the device still answers
    rc, _, _ = _call_noalloc(L, s, g, idx, h, hn, u)
    assert rc == 0
