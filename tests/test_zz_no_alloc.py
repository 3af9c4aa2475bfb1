"""nmfb_gram_static: the Gram matrix of a factor without a caller workspace (K-split slots in a device array built
into the library; include/nmf_b200.h).  No counterpart in the reference -- the entry point exists for callers that
cannot allocate device memory.  Checked through the C ABI against nmfb_gram (caller workspace) on two calls in a row
with different factors (the array keeps the previous call's contents), against float64 NumPy, and for the rejection
of a shape whose slots exceed the array.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-4          # the bar of every GPU parity test here; the two entry points run the same kernels (expect 0)


def _factor(lib, torch, dev, rng, L, k, KP):
    """Device buffers of one factor in the library's layout, filled through nmfb_import_factor."""
    from nmf_b200._lib import Factor, check
    Lp = (L + 255) // 256 * 256
    bufs = dict(master=torch.zeros((KP, Lp), dtype=torch.float32, device=dev),
                cplanes=torch.zeros((3, 256, Lp), dtype=torch.bfloat16, device=dev),
                rplanes=torch.zeros((2, Lp, KP), dtype=torch.bfloat16, device=dev))
    fs = Factor(bufs["master"].data_ptr(), bufs["cplanes"].data_ptr(), bufs["rplanes"].data_ptr(), L, Lp, 0, 0)
    host = rng.random((L, k)).astype(np.float32)                      # (L, k), like W
    src = torch.from_numpy(host).to(dev)
    check(lib.nmfb_import_factor(src.data_ptr(), 0, k, 1, k, KP, C.byref(fs), None))
    torch.cuda.synchronize(dev)
    return fs, bufs, host


@pytest.mark.parametrize("k,KP,nplanes", [(100, 128, 2), (48, 64, 3)])
def test_gram_static_matches_gram_on_two_calls_in_a_row(k, KP, nplanes):
    import torch
    from nmf_b200._lib import check, load
    lib = load()
    dev = torch.device("cuda:0")
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    rng = np.random.default_rng(5)
    assert int(lib.nmfb_gram_static_max_bytes()) == 8 << 20
    with torch.cuda.device(dev):
        cases = [_factor(lib, torch, dev, rng, L, k, KP) for L in (20000, 5000)]   # second call: fewer, other slots
        pbytes = max(int(lib.nmfb_gram_partial_bytes(k, L, sms)) for L in (20000, 5000))
        partial = torch.empty(pbytes // 4, dtype=torch.float32, device=dev)
        out = torch.zeros((2, 2, KP, KP), dtype=torch.float32, device=dev)         # [call][entry point]
        out64 = torch.zeros((2, 2, KP, KP), dtype=torch.float64, device=dev)
        # one untimed call of each entry point first: the driver loads a module's code and data (the built-in array
        # included) when it is first touched, which moves the free-memory figure the check below reads
        warm = torch.zeros((KP, KP), dtype=torch.float32, device=dev), torch.zeros((KP, KP), dtype=torch.float64,
                                                                                   device=dev)
        check(lib.nmfb_gram(C.byref(cases[1][0]), KP, nplanes, partial.data_ptr(), pbytes, warm[0].data_ptr(),
                            warm[1].data_ptr(), sms, None))
        check(lib.nmfb_gram_static(C.byref(cases[1][0]), KP, nplanes, warm[0].data_ptr(), warm[1].data_ptr(), sms,
                                   None))
        torch.cuda.synchronize(dev)
        free_before = torch.cuda.mem_get_info(dev)[0]
        for call, (fs, _bufs, _host) in enumerate(cases):
            check(lib.nmfb_gram_static(C.byref(fs), KP, nplanes, out[call, 0].data_ptr(), out64[call, 0].data_ptr(),
                                       sms, None))
        torch.cuda.synchronize(dev)
        assert torch.cuda.mem_get_info(dev)[0] == free_before                      # nothing was allocated
        for call, (fs, _bufs, _host) in enumerate(cases):
            check(lib.nmfb_gram(C.byref(fs), KP, nplanes, partial.data_ptr(), pbytes, out[call, 1].data_ptr(),
                                out64[call, 1].data_ptr(), sms, None))
        torch.cuda.synchronize(dev)
    for call, (_fs, _bufs, host) in enumerate(cases):
        a64, b64 = out64[call, 0].cpu().numpy(), out64[call, 1].cpu().numpy()
        a32, b32 = out[call, 0].cpu().numpy(), out[call, 1].cpu().numpy()
        ref = host.astype(np.float64).T @ host.astype(np.float64)
        err_ab = np.linalg.norm(a64 - b64) / np.linalg.norm(b64)
        err_ref = np.linalg.norm(a64[:k, :k] - ref) / np.linalg.norm(ref)
        print(f"k={k} call {call}: static vs caller-workspace {err_ab:.2e}, static vs float64 NumPy {err_ref:.2e}")
        assert err_ab < TOL and np.linalg.norm(a32 - b32) / np.linalg.norm(b32) < TOL
        assert err_ref < TOL
        assert np.all(a64[k:, :] == 0) and np.all(a64[:, k:] == 0)                 # padded components stay zero


def test_gram_static_rejects_a_shape_beyond_its_scratch():
    import torch
    from nmf_b200._lib import Factor, load
    lib = load()
    dev = torch.device("cuda:0")
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    KP, L = 256, 262144                                # > 50 K-splits x 256 x 256 x 4 bytes of slots
    with torch.cuda.device(dev):
        g32 = torch.zeros((KP, KP), dtype=torch.float32, device=dev)
        g64 = torch.zeros((KP, KP), dtype=torch.float64, device=dev)
        fs = Factor(g32.data_ptr(), g32.data_ptr(), g32.data_ptr(), L, L, 0, 0)     # never dereferenced: plan stage
        rc = lib.nmfb_gram_static(C.byref(fs), KP, 2, g32.data_ptr(), g64.data_ptr(), sms, None)
    msg = lib.nmfb_last_error().decode()
    assert rc == -1, (rc, msg)                         # NMFB_ERR_ARG
    assert "built-in scratch" in msg and str(8 << 20) in msg
