"""b2seg_seg_metrics_from_probs_noalloc (static device scratch, no cudaMalloc) against the
workspace entry point b2seg_seg_metrics_from_probs, through the C ABI.

Collected last on purpose.  The entry point has no counterpart
in the reference (include/b2seg.h).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

COUNTS = (6, 7, 8, 10, 11, 12, 13)   # S_PB, S_PB_T, S_N, S_TP, S_FP, S_FN, S_TN: exact integers
SUMS = (0, 1, 2)                     # sum p, sum t, sum p*t


def _inputs(n, seed, dev, offset=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    p = torch.rand(n + offset, device=dev, generator=g)[offset:]
    t = (torch.rand(n + offset, device=dev, generator=g) > 0.8).float()[offset:]
    return p, t


def _workspace_variant(L, p, t):
    ws = torch.zeros(L.lib().b2seg_seg_reduce_workspace_bytes(), dtype=torch.uint8, device=p.device)
    out = torch.full((L.STATS_LEN,), -1.0, dtype=torch.float64, device=p.device)
    L.check(L.lib().b2seg_seg_metrics_from_probs(p.data_ptr(), t.data_ptr(), p.numel(), ws.data_ptr(),
                                                 out.data_ptr(), L.stream_ptr()))
    return out.cpu().numpy()


def _static_variant(L, p, t, out):
    out.fill_(-1.0)
    L.check(L.lib().b2seg_seg_metrics_from_probs_noalloc(p.data_ptr(), t.data_ptr(), p.numel(),
                                                         out.data_ptr(), L.stream_ptr()))
    return out.cpu().numpy()


def _assert_same(got, want):
    for i in COUNTS:
        assert got[i] == want[i], (i, got[i], want[i])
    for i in SUMS:
        assert got[i] == pytest.approx(want[i], rel=1e-12), (i, got[i], want[i])


def test_static_scratch_matches_workspace_variant_on_consecutive_calls(cuda):
    from b2seg import _lib as L
    out = torch.empty(L.STATS_LEN, dtype=torch.float64, device=cuda)
    # two calls in a row on different data and sizes (the second sees the first's partial rows),
    # then a small one-block call and a misaligned one (scalar-load kernel)
    cases = [(4 * 256 * 256, 11, 0), (64 * 256 * 256, 12, 0), (1000, 13, 0), (3 * 100 * 100 + 3, 14, 1)]
    for n, seed, off in cases:
        p, t = _inputs(n, seed, cuda, off)
        want = _workspace_variant(L, p, t)
        got = _static_variant(L, p, t, out)
        _assert_same(got, want)
        # counts against torch on the same data
        pb = (p > 0.5).double()
        tp = float((pb * t.double()).sum())
        assert got[10] == tp and got[11] == float(pb.sum()) - tp and got[8] == n
    # same input twice in a row: identical bits (fixed reduction order, stale scratch is harmless)
    p, t = _inputs(16 * 256 * 256, 15, cuda)
    a = _static_variant(L, p, t, out).copy()
    b = _static_variant(L, p, t, out).copy()
    assert np.array_equal(a, b)
    _assert_same(a, _workspace_variant(L, p, t))


def test_static_scratch_call_does_not_allocate_device_memory(cuda):
    from b2seg import _lib as L
    p, t = _inputs(8 * 256 * 256, 21, cuda)
    out = torch.empty(L.STATS_LEN, dtype=torch.float64, device=cuda)
    _static_variant(L, p, t, out)           # first call on the device: symbol lookup, module load
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    for _ in range(3):
        L.check(L.lib().b2seg_seg_metrics_from_probs_noalloc(p.data_ptr(), t.data_ptr(), p.numel(),
                                                             out.data_ptr(), L.stream_ptr()))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_reserved() == reserved0
    assert free1 >= free0, (free0, free1)


def test_static_scratch_rejects_sizes_beyond_its_maximum(cuda):
    from b2seg import _lib as L
    import re, os
    hdr = open(os.path.join(os.path.dirname(__file__), "..", "include", "b2seg.h")).read()
    shift = int(re.search(r"#define\s+B2SEG_NOALLOC_MAX_PIXELS\s+\(1LL\s*<<\s*(\d+)\)", hdr).group(1))
    p, t = _inputs(1024, 31, cuda)
    out = torch.zeros(L.STATS_LEN, dtype=torch.float64, device=cuda)
    rc = L.lib().b2seg_seg_metrics_from_probs_noalloc(p.data_ptr(), t.data_ptr(), (1 << shift) + 1,
                                                      out.data_ptr(), L.stream_ptr())
    assert rc != 0
    msg = L.lib().b2seg_last_error().decode()
    assert "maximum" in msg and "noalloc" in msg, msg
    assert L.lib().b2seg_seg_metrics_from_probs_noalloc(p.data_ptr(), t.data_ptr(), 0, out.data_ptr(),
                                                        L.stream_ptr()) != 0
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0.0    # nothing was launched
