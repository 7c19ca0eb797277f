"""GPU test of the allocation-free entry point mc_eval_batch_colex32_dev_noalloc (include/mcpoker.h): its class ids must
equal those of the existing entry point mc_eval_batch_colex32 bit for bit (integer work: no tolerance), on two calls in a
row that share the library's static scratch array.  Collected last on purpose (file name):
a failure here must not hide the other tests.  Passes alone in a fresh process and after the other tests
(the `gpu` fixture initialises the engine either way)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_call(E, torch, ranks):
    """ranks (numpy uint32) -> class ids (numpy uint16) through the C ABI on device buffers of the caller"""
    L = E.lib()
    n = int(ranks.shape[0])
    d_ranks = torch.from_numpy(ranks.view(np.int32).copy()).cuda()
    d_cls = torch.full((max(n, 2),), -1, dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    E.check(L.mc_eval_batch_colex32_dev_noalloc(d_ranks.data_ptr(), n, d_cls.data_ptr(), stream))
    torch.cuda.synchronize()
    return d_cls[:n].cpu().numpy().view(np.uint16)


def test_noalloc_entry_point_equals_colex32_on_two_calls_in_a_row(gpu):
    import torch
    E, api = gpu.engine, gpu.api
    torch.cuda.set_device(0)                                   # the engine drives device 0 (and more on a multi-GPU box)
    rng = np.random.default_rng(20261105)
    total = 133784560
    # first call: the maximum size, with a few ranks past the last combination (-> MC_CLASS_INVALID)
    a = rng.integers(0, total, E.NOALLOC_MAX_HANDS, dtype=np.uint32)
    a[:3] = (total, 0xFFFFFFFF, total - 1)
    # second call: other hands, smaller and odd-sized, so the array still holds the first call's masks behind them
    b = rng.integers(0, total, 300_001, dtype=np.uint32)
    for ranks in (a, b):
        got = _device_call(E, torch, ranks)
        want = api.eval_batch_colex32(ranks)
        assert np.array_equal(got, want)
    assert got.shape == (300_001,) and int(_device_call(E, torch, a)[0]) == E.CLASS_INVALID
    # sizes: nothing to do, one hand, one more than the scratch holds
    assert _device_call(E, torch, np.zeros(0, dtype=np.uint32)).shape == (0,)
    assert np.array_equal(_device_call(E, torch, b[:1]), api.eval_batch_colex32(b[:1]))
    L = E.lib()
    d = torch.zeros(8, dtype=torch.int32, device="cuda")
    rc = L.mc_eval_batch_colex32_dev_noalloc(d.data_ptr(), E.NOALLOC_MAX_HANDS + 1, d.data_ptr(), None)
    assert rc == E.MC_ERR_INVALID and b"MC_NOALLOC_MAX_HANDS" in L.mc_last_error()
    assert L.mc_eval_batch_colex32_dev_noalloc(None, 4, d.data_ptr(), None) == E.MC_ERR_INVALID


def test_noalloc_entry_point_allocates_nothing(gpu):
    """free device memory is the same before and after a call (the buffers are the caller's, made before the measurement)"""
    import torch
    E = gpu.engine
    L = E.lib()
    torch.cuda.set_device(0)
    n = 1 << 18
    d_ranks = torch.arange(n, dtype=torch.int32, device="cuda")
    d_cls = torch.empty(n, dtype=torch.int16, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    E.check(L.mc_eval_batch_colex32_dev_noalloc(d_ranks.data_ptr(), n, d_cls.data_ptr(), stream))     # (first use: module state settled)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        E.check(L.mc_eval_batch_colex32_dev_noalloc(d_ranks.data_ptr(), n, d_cls.data_ptr(), stream))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0
    # ranks 0..n-1 are the first hands of the colex enumeration: class ids equal to the host-buffer call's
    want = gpu.api.eval_batch_colex32(np.arange(n, dtype=np.uint32))
    assert np.array_equal(d_cls.cpu().numpy().view(np.uint16), want)
