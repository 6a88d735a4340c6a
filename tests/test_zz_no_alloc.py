"""Allocation-free batched SYMV (``gadmm_symv_batched_noalloc``), exercised through the C ABI.

Some users of the native library cannot allocate device memory when they need ``y = S x``: code
loaded into an application that claimed the whole device for its own arena at start-up, or a
control loop that must never hit the device-wide synchronisation of cudaMalloc/cudaFree.  For
them the library carries a small fixed work area of its own.  This test drives that entry point
exactly as such a user would (raw pointers, ctypes, no torch wrapper around the call) and checks
it against the ordinary entry point, whose scratch the caller supplies.

Two calls are made back to back on purpose: the built-in work area is shared by all calls and
still holds the first call's partial sums and tickets when the second one starts, so the second
result is only right if the entry point re-initialises what it must.

Collected last (file name) so that a ``-x`` run has already executed every other test.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _packed_problem(N, d, seed, store, dev):
    from gadmm_b200.ops import linalg as LA
    from gadmm_b200.ops import native as nv
    dp = nv.padded(d)
    g = torch.Generator(device="cpu").manual_seed(seed)
    S = torch.randn(N, dp, dp, dtype=torch.float64, generator=g).to(dev)
    S = S + S.transpose(1, 2)
    P = LA.pack_sym(S, d, store)
    ptrs = LA.ptr_table(P)
    xs = [torch.randn(N, dp, dtype=torch.float64, generator=g).to(dev) for _ in range(2)]
    return dp, P, ptrs, xs


@pytest.mark.parametrize("store_name", ["f64", "f32"])
def test_noalloc_symv_matches_regular_entry_point_twice(store_name):
    from gadmm_b200.ops import native as nv
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    L = nv.lib()
    L.gadmm_symv_batched_noalloc.restype = C.c_int
    L.gadmm_symv_batched.restype = C.c_int
    store = nv.STORE_F64 if store_name == "f64" else nv.STORE_F32
    N, d = 5, 300                      # nT = 5: odd tile-row count, several tiles per matrix
    dp, P, ptrs, xs = _packed_problem(N, d, 11, store, dev)
    assert N * nv.symv_scratch_elems(dp) <= L.gadmm_symv_noalloc_max_scratch_elems()

    # ordinary entry point: the caller brings scratch and zeroed counters
    scratch = torch.empty(N * nv.symv_scratch_elems(dp), dtype=torch.float64, device=dev)
    counters = torch.zeros(N, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    want, got = [], []
    for x in xs:
        y = torch.empty_like(x)
        rc = L.gadmm_symv_batched(C.c_void_p(ptrs.data_ptr()), C.c_void_p(x.data_ptr()),
                                  C.c_void_p(y.data_ptr()), N, dp, store,
                                  C.c_void_p(scratch.data_ptr()), C.c_void_p(counters.data_ptr()),
                                  C.c_void_p(stream))
        assert rc == 0, L.gadmm_last_error()
        want.append(y)
    # allocation-free entry point: inputs and outputs only, two calls in a row
    outs = [torch.empty_like(xs[0]) for _ in xs]     # outputs exist before the calls
    torch.cuda.synchronize()
    free_before = None
    for i, (x, y) in enumerate(zip(xs, outs)):
        if i == 1:                                   # set-up happened in call 0
            free_before = torch.cuda.mem_get_info(dev)[0]
        rc = L.gadmm_symv_batched_noalloc(C.c_void_p(ptrs.data_ptr()), C.c_void_p(x.data_ptr()),
                                          C.c_void_p(y.data_ptr()), N, dp, store,
                                          C.c_void_p(stream))
        assert rc == 0, L.gadmm_last_error()
        got.append(y)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] >= free_before      # the second call took no memory
    for w, g_ in zip(want, got):
        torch.testing.assert_close(g_, w, rtol=1e-12, atol=1e-9)
    assert not torch.equal(got[0], got[1])           # the two calls really were different problems


def test_noalloc_symv_rejects_oversized_requests():
    from gadmm_b200.ops import native as nv
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    L = nv.lib()
    L.gadmm_symv_batched_noalloc.restype = C.c_int
    L.gadmm_symv_noalloc_max_scratch_elems.restype = C.c_size_t
    cap = L.gadmm_symv_noalloc_max_scratch_elems()
    dp = 4096
    N = cap // nv.symv_scratch_elems(dp) + 1          # one matrix too many for the built-in area
    dummy = torch.zeros(8, dtype=torch.float64, device=dev)
    p = C.c_void_p(dummy.data_ptr())
    rc = L.gadmm_symv_batched_noalloc(p, p, p, int(N), dp, nv.STORE_F64, C.c_void_p(0))
    assert rc != 0
    msg = L.gadmm_last_error().decode()
    assert "scratch" in msg and str(cap) in msg
    rc = L.gadmm_symv_batched_noalloc(p, p, p, L.gadmm_symv_noalloc_max_matrices() + 1, 64,
                                      nv.STORE_F64, C.c_void_p(0))
    assert rc != 0 and "limit" in L.gadmm_last_error().decode()
    torch.cuda.synchronize()                          # nothing was launched, nothing faulted
