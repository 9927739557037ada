"""somar_norm_noalloc (the allocation-free entry point added; no counterpart in
the reference) against the existing entry point somar_norm, through the C ABI.

Named test_zz_* so that pytest collects it LAST: a failure here must not hide
the parity tests.  Passes alone in a fresh process and after the other tests (the static scratch array keeps the last
call's contents; the entry point clears what has to start at zero).
"""
import numpy as np
import pytest

from common import Problem

pytestmark = pytest.mark.gpu


def _level_and_field(somar, n, seed):
    pb = Problem(n, map_name="stretched", aniso=4.0, seed=seed)
    lev = pb.level(somar)
    x = pb.rand()
    return lev, lev.field(0, x), x


@pytest.mark.parametrize("n", [(48, 40, 24), (130, 33, 17)])
def test_norm_noalloc_matches_somar_norm_twice(n, somar):
    import torch
    dev = torch.device("cuda", 0)
    out = torch.full((1,), -7.0, dtype=torch.float64, device=dev)   # the caller supplies the output
    for call in range(2):                                            # two calls in a row share the static scratch
        lev, f, x = _level_and_field(somar, n, seed=11 + call)
        xd = torch.from_numpy(np.ascontiguousarray(x)).to(dev)      # the caller supplies the input
        torch.cuda.synchronize()
        for ord_ in (0, 1, 2):
            want = lev.norm(f, ord_)                                 # existing entry point (k_reduce, level scratch)
            somar.norm_noalloc(xd, ord_, out)                        # new entry point (static __device__ scratch)
            torch.cuda.synchronize()
            got = float(out.item())
            if ord_ == 0:
                assert got == want, (call, got, want)                # a maximum is order-independent: bit-exact
            else:
                assert abs(got - want) <= 1e-13 * abs(want), (call, ord_, got, want)   # the operator-parity tolerance
        lev.destroy()


def test_norm_noalloc_does_not_allocate_and_rejects_oversize(somar):
    import torch
    dev = torch.device("cuda", 0)
    x = torch.linspace(-3.0, 2.0, 100001, dtype=torch.float64, device=dev)
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    somar.norm_noalloc(x, 0, out)          # first call on the device: takes the symbol's address; the driver may also
    torch.cuda.synchronize()               # load the kernel's code lazily here, so the memory baseline is taken after it
    free0, _ = torch.cuda.mem_get_info(dev)
    for _ in range(20):
        somar.norm_noalloc(x, 2, out)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 == free0, (free0, free1)   # device memory untouched by 20 calls
    assert abs(float(out.item()) - float(torch.linalg.vector_norm(x).item())) <= 1e-12 * float(out.item())
    # beyond the stated maximum: a clear error, not a fallback allocation (no memory of that size is touched: the
    # size check comes first, so a small tensor with a large claimed n is enough)
    import ctypes as C
    rc = somar.lib().somar_norm_noalloc(C.c_void_p(x.data_ptr()), C.c_longlong(somar.NOALLOC_MAX_N + 1), 0,
                                        C.c_void_p(out.data_ptr()), C.c_void_p(0))
    assert rc == -1 and b"SOMAR_NOALLOC_MAX_N" in somar.lib().somar_last_error()
    torch.cuda.synchronize()
