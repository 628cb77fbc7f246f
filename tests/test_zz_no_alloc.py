"""orbx_knn2_device_static: the allocation-free 2-NN entry point (file-scope device scratch, include/orbx.h) against
orbx_knn2 on the same descriptors -- bit-exact indices, distances and ratio flags, on calls in a row that share the scratch.
Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAXQ = 16384   # ORBX_KNN2_STATIC_MAX_QUERIES


def _static_knn2(lib, torch, q, t, stream=None, want_good=True, check_mem=True):
    nq, nt = len(q), len(t)
    dq = torch.from_numpy(q).cuda() if nq else torch.empty((0, 32), dtype=torch.uint8, device="cuda")
    dt = torch.from_numpy(t).cuda() if nt else torch.empty((0, 32), dtype=torch.uint8, device="cuda")
    # poisoned outputs: every element must be written by the call
    didx = torch.full((max(nq, 1), 2), -7, dtype=torch.int32, device="cuda")
    ddist = torch.full((max(nq, 1), 2), -7, dtype=torch.int32, device="cuda")
    dgood = torch.full((max(nq, 1),), 9, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = lib.orbx_knn2_device_static(C.c_void_p(dq.data_ptr() if nq else None), nq, C.c_void_p(dt.data_ptr() if nt else None), nt,
                                     C.c_void_p(didx.data_ptr()), C.c_void_p(ddist.data_ptr()),
                                     C.c_void_p(dgood.data_ptr() if want_good else None), C.c_void_p(stream))
    assert rc == 0, lib.orbx_last_error()
    torch.cuda.synchronize()
    if check_mem:   # the call took no device memory (the first launch of a kernel may still load its code lazily)
        assert torch.cuda.mem_get_info()[0] >= free0
    return didx.cpu().numpy()[:nq], ddist.cpu().numpy()[:nq], dgood.cpu().numpy()[:nq]


def test_static_knn2_equals_knn2_on_calls_in_a_row(orbx, synth):
    import torch
    lib = orbx.lib()
    m = orbx.ORBmatcher()
    try:
        q1, t1 = synth.random_descriptors(3000, 11), synth.random_descriptors(5000, 12)
        t1[77] = t1[5]; q1[9] = t1[5]; q1[:400, 6:] = 0   # exact ties: the lower train index wins
        t1b = t1.copy(); t1b[:, 6:] = 0
        # second call: fewer queries, other data -- stale keys of the first call must not show
        q2, t2 = synth.random_descriptors(1234, 21), synth.random_descriptors(777, 22)
        cases = [(q1, t1), (q2, t2), (q1[:400], t1b), (q2, t2[:1]), (q2[:50], t2[:0]), (q1, t1)]
        st = torch.cuda.Stream()
        for n, (q, t) in enumerate(cases):
            want = m.knn2(q, t)
            got = _static_knn2(lib, torch, q, t, stream=st.cuda_stream if n % 2 else None, check_mem=n > 0)
            for w, g, name in zip(want, got, ("idx", "dist", "good")):
                assert np.array_equal(w, g), (n, name)
        # good is optional
        got = _static_knn2(lib, torch, q2, t2, want_good=False)
        assert np.array_equal(got[0], m.knn2(q2, t2)[0]) and (got[2] == 9).all()
    finally:
        m.close()


def test_static_knn2_limit_and_arguments(orbx, synth):
    import torch
    lib = orbx.lib()
    m = orbx.ORBmatcher()
    try:
        t = synth.random_descriptors(300, 5)
        q = synth.random_descriptors(MAXQ, 6)
        want = m.knn2(q, t)
        got = _static_knn2(lib, torch, q, t, check_mem=False)   # exactly the maximum is served
        assert all(np.array_equal(w, g) for w, g in zip(want, got))
        dq = torch.zeros((MAXQ + 1, 32), dtype=torch.uint8, device="cuda")
        out = torch.zeros((MAXQ + 1, 2), dtype=torch.int32, device="cuda")
        p = lambda x: C.c_void_p(x.data_ptr())
        rc = lib.orbx_knn2_device_static(p(dq), MAXQ + 1, p(dq), 10, p(out), p(out), None, None)
        assert rc == orbx.ORBX_ERR_CAPACITY and b"ORBX_KNN2_STATIC_MAX_QUERIES" in lib.orbx_last_error()
        assert lib.orbx_knn2_device_static(None, 5, p(dq), 10, p(out), p(out), None, None) == orbx.ORBX_ERR_ARG
        assert lib.orbx_knn2_device_static(p(dq), -1, p(dq), 10, p(out), p(out), None, None) == orbx.ORBX_ERR_ARG
        assert lib.orbx_knn2_device_static(p(dq), 0, None, 0, None, None, None, None) == 0
    finally:
        m.close()
