"""b200sift_match_device_noalloc (include/b200sift.h): the matcher with its scratch in a static device array, for callers
that cannot call cudaMalloc. No counterpart in the reference.

Through the C ABI (ctypes): its output must equal b200sift_match_device's -- match indices, i.e. the bit-exact bar every
matcher test in this repository uses -- on two calls in a row with different sizes and data (the array keeps the previous
call's contents: the second call must not see them), sizes beyond the stated maximum are rejected with a clear error, and
a call leaves the device's free memory unchanged.

Collected last on purpose (file name): a failure here must not hide other tests. It has to
pass alone in a fresh process (the new entry point is called BEFORE the existing one, so its one-time set-up runs on a
device nothing else has touched) and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _lib():
    from another_cuda_sift_b200 import _lib

    return _lib, _lib.load()


def _dev(a):
    import torch

    return torch.from_numpy(np.ascontiguousarray(a).view(np.int16)).cuda()


def test_noalloc_matcher_equals_matcher_two_calls_in_a_row():
    import torch

    from another_cuda_sift_b200.synth import synth_descriptors

    mod, lib = _lib()
    max_rows = lib.b200sift_match_noalloc_max_rows()
    assert max_rows == 8192
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    # two problems, different sizes and data; the second is smaller in both dimensions, so stale survivor records and
    # norms of the first call lie all around what it uses
    q1, t1 = synth_descriptors(2000, 2000)
    rng = np.random.default_rng(11)
    q2, t2 = synth_descriptors(1500, 5000)
    q2, t2 = q2[rng.permutation(1500)[:700]], t2[rng.permutation(5000)[:4100]]
    cases = [(q1, t1), (q2, t2)]
    bufs = [(_dev(q), _dev(t), torch.full((len(q),), -7, dtype=torch.int32, device="cuda"),
             torch.full((len(q),), -9, dtype=torch.int32, device="cuda")) for q, t in cases]
    torch.cuda.synchronize()

    got = []
    for (q, t), (dq, dt, out_new, _) in zip(cases, bufs):   # the new entry point FIRST, two calls back to back
        mod.check(lib.b200sift_match_device_noalloc(dq.data_ptr(), len(q), dt.data_ptr(), len(t), out_new.data_ptr(), stream))
    torch.cuda.synchronize()
    for (q, t), (dq, dt, out_new, out_ref) in zip(cases, bufs):              # then the existing entry point
        mod.check(lib.b200sift_match_device(dq.data_ptr(), len(q), dt.data_ptr(), len(t), out_ref.data_ptr(), stream))
        torch.cuda.synchronize()
        a, b = out_new.cpu().numpy(), out_ref.cpu().numpy()
        assert np.array_equal(a, b)
        assert (b >= 0).sum() > 0.1 * len(b) and (b >= -1).all() and (b < len(t)).all()
        got.append(b)
    # once more in the opposite order of sizes: results are reproduced whatever the array held before, and -- every
    # kernel involved having run once (lazy module loading is done) -- a call leaves the device's free memory untouched
    for k in (1, 0, 1):
        (q, t), (dq, dt, out_new, _) = cases[k], bufs[k]
        out_new.fill_(-7)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        mod.check(lib.b200sift_match_device_noalloc(dq.data_ptr(), len(q), dt.data_ptr(), len(t), out_new.data_ptr(), stream))
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before, "the no-alloc entry point changed the device's free memory"
        assert np.array_equal(out_new.cpu().numpy(), got[k])


def test_noalloc_matcher_edges_and_limit():
    import torch

    from another_cuda_sift_b200.synth import synth_descriptors

    mod, lib = _lib()
    max_rows = lib.b200sift_match_noalloc_max_rows()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    q, t = synth_descriptors(300, max_rows)
    dq, dt = _dev(q), _dev(t)
    out_new = torch.full((300,), -7, dtype=torch.int32, device="cuda")
    out_ref = torch.full((300,), -9, dtype=torch.int32, device="cuda")
    # exactly at the limit: accepted, equal to the existing entry point
    mod.check(lib.b200sift_match_device_noalloc(dq.data_ptr(), 300, dt.data_ptr(), max_rows, out_new.data_ptr(), stream))
    mod.check(lib.b200sift_match_device(dq.data_ptr(), 300, dt.data_ptr(), max_rows, out_ref.data_ptr(), stream))
    torch.cuda.synchronize()
    assert np.array_equal(out_new.cpu().numpy(), out_ref.cpu().numpy())
    # no train rows: -1 everywhere (Match.cu:125-129 sentinels), nq == 0: nothing to do
    mod.check(lib.b200sift_match_device_noalloc(dq.data_ptr(), 300, dt.data_ptr(), 0, out_new.data_ptr(), stream))
    torch.cuda.synchronize()
    assert bool((out_new == -1).all())
    assert lib.b200sift_match_device_noalloc(dq.data_ptr(), 0, dt.data_ptr(), 10, out_new.data_ptr(), stream) == 0
    # beyond the limit: rejected, nothing launched, the message names the limit
    for nq, nt in ((300, max_rows + 1), (max_rows + 1, 300)):
        rc = lib.b200sift_match_device_noalloc(dq.data_ptr(), nq, dt.data_ptr(), nt, out_new.data_ptr(), stream)
        assert rc == -1  # B200SIFT_EINVAL
        msg = lib.b200sift_last_error().decode()
        assert str(max_rows) in msg and "noalloc" in msg, msg
    with pytest.raises(mod.B200SiftError):
        mod.check(lib.b200sift_match_device_noalloc(None, 5, dt.data_ptr(), 5, out_new.data_ptr(), stream))
