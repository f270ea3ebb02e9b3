"""ab_confusion_per_pixel_noalloc (include/annonet_b200.h; annonet_b200/csrc/noalloc.cu): the per-pixel confusion matrix on
caller-owned device buffers with the counters in the library's static __device__ scratch -- no allocation inside the call.
The entry point has no counterpart in the reference; it is checked against the
existing entry point, ab_evaluate_labels, whose kernel it is a variant of.  Integer counts: the bar is bit-exact.

Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _labels(rng, h, w, k, ignore_fraction=0.1, stray=True):
    t = rng.integers(0, k, size=(h, w)).astype(np.uint16)
    p = np.where(rng.random((h, w)) < 0.7, t, rng.integers(0, k, size=(h, w))).astype(np.uint16)
    t[rng.random((h, w)) < ignore_fraction] = 65535          # not labelled: drops out of the matrix
    if stray:
        p[rng.random((h, w)) < 0.01] = k + 3                 # a prediction outside [0, K): drops out too
    return t, p


def _noalloc(ab, torch, truth, pred, k, out=None, stream=None):
    """through the C ABI, with torch only as the CALLER's allocator of inputs and output"""
    dev = torch.device("cuda", 0)
    # uint16 has no arithmetic in torch: ship the bytes as int16, the kernel reads them as uint16
    td = torch.from_numpy(truth.view(np.int16).copy()).to(dev)
    pd = torch.from_numpy(pred.view(np.int16).copy()).to(dev)
    if out is None:
        out = torch.full((k * k,), -1, dtype=torch.int64, device=dev)     # poisoned: every cell must be written
    st = torch.cuda.current_stream(dev) if stream is None else stream
    torch.cuda.synchronize(dev)
    before = torch.cuda.memory_allocated(dev)
    ab.confusion_per_pixel_noalloc(td.data_ptr(), pd.data_ptr(), truth.size, k, out.data_ptr(), st.cuda_stream)
    assert torch.cuda.memory_allocated(dev) == before          # (torch's own allocator was not touched by the call)
    st.synchronize()
    return out.cpu().numpy().astype(np.uint64).reshape(k, k), out


def test_noalloc_confusion_matches_the_existing_entry_point_twice_in_a_row(ab):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(2024)
    net = ab.RuntimeNet(0)
    # two calls in a row on DIFFERENT inputs and class counts: the static scratch keeps the first call's counts, the second
    # call has to clear what it uses (a larger matrix first, so that stale cells lie inside and outside the second one)
    cases = [(_labels(rng, 517, 389, 11), 11), (_labels(rng, 300, 420, 5), 5), (_labels(rng, 64, 64, 64, stray=False), 64)]
    for (truth, pred), k in cases:
        want, _ = net.evaluate_labels(truth, pred, k)           # the existing entry point (per-pixel matrix)
        got, _ = _noalloc(ab, torch, truth, pred, k)
        assert got.dtype == want.dtype and np.array_equal(got, want), f"K = {k}"
        labelled = (truth < k) & (pred < k)
        assert int(got.sum()) == int(labelled.sum())
    # the same input again, into the same output buffer, on a stream of the caller's: identical (nothing accumulates across calls)
    (truth, pred), k = cases[0]
    want, _ = net.evaluate_labels(truth, pred, k)
    side = torch.cuda.Stream(device=0)
    got1, buf = _noalloc(ab, torch, truth, pred, k, stream=side)
    got2, _ = _noalloc(ab, torch, truth, pred, k, out=buf, stream=side)
    assert np.array_equal(got1, want) and np.array_equal(got2, want)


def test_noalloc_edge_cases_and_limit(ab):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(7)
    # empty input: a zero matrix (the scratch still holds the previous call's counts when this runs)
    out = torch.full((9,), -1, dtype=torch.int64, device="cuda:0")
    ab.confusion_per_pixel_noalloc(0, 0, 0, 3, out.data_ptr(), 0)
    torch.cuda.synchronize()
    assert not out.cpu().numpy().any()
    # one class
    truth, pred = _labels(rng, 33, 77, 1)
    want, _ = ab.RuntimeNet(0).evaluate_labels(truth, pred, 1)
    got, _ = _noalloc(ab, torch, truth, pred, 1)
    assert np.array_equal(got, want)
    # beyond the scratch: rejected with the limit in the message, nothing launched
    assert ab.NOALLOC_MAX_CLASSES == 64
    with pytest.raises(ValueError, match="64"):
        ab.confusion_per_pixel_noalloc(out.data_ptr(), out.data_ptr(), 4, ab.NOALLOC_MAX_CLASSES + 1, out.data_ptr(), 0)
    with pytest.raises(ValueError):
        ab.confusion_per_pixel_noalloc(out.data_ptr(), out.data_ptr(), 4, 0, out.data_ptr(), 0)
    with pytest.raises(ValueError):
        ab.confusion_per_pixel_noalloc(out.data_ptr(), out.data_ptr(), 4, 3, 0, 0)
    assert ab.lib.ab_confusion_per_pixel_noalloc(None, None, C.c_longlong(-1), 3, C.c_void_p(out.data_ptr()), None) == 1   # AB_ERR_ARGUMENT
    torch.cuda.synchronize()
