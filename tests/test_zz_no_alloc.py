"""odvae_colsum_static: the allocation-free variant of odvae_colsum (include/odvae_b200.h; no counterpart in the reference,
added for callers that cannot call cudaMalloc). Its stage-1 partial sums live in a file-scope `static __device__` array
inside the library, looked up with cudaGetSymbolAddress once per device. Called through the C ABI (ctypes, odvae_b200/lib.py)
and compared with the existing workspace entry point on two calls in a row (the second with a different shape, so the array
still holds the first call's partials). This file is collected last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_COLS = 512   # ODVAE_COLSUM_STATIC_MAX_COLS
TOL = 2e-4       # tests/test_kernels_gpu.py::test_colsum: |out - ref| < TOL * max column sum of |m|


def _both(lib, m, rows, cols, ld, code, accumulate=0, init=None):
    dev = m.device
    out_ws = torch.zeros(cols, device=dev) if init is None else init.clone()
    out_st = torch.zeros(cols, device=dev) if init is None else init.clone()
    nbytes = lib.call("colsum_workspace_bytes", cols)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    lib.call("colsum", m, rows, cols, ld, code, out_ws, accumulate, ws, ws.numel())
    torch.cuda.synchronize()
    # inputs, outputs and everything torch allocates exist before the call: the entry point itself must allocate nothing
    alloc0, free0 = torch.cuda.memory_allocated(dev), torch.cuda.mem_get_info(dev)[0]
    lib.call("colsum_static", m, rows, cols, ld, code, out_st, accumulate)
    torch.cuda.synchronize()
    alloc1, free1 = torch.cuda.memory_allocated(dev), torch.cuda.mem_get_info(dev)[0]
    return out_ws, out_st, (alloc0, free0, alloc1, free1)


def test_colsum_static_matches_colsum_on_two_calls_in_a_row():
    from odvae_b200 import lib
    lib.require_device()
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(5)
    warm = torch.randn(64, 64, generator=g).to(dev)
    lib.call("colsum_static", warm, 64, 64, 64, lib.F32, torch.empty(64, device=dev), 0)  # first use: symbol address lookup
    torch.cuda.synchronize()
    # call 1: the widest shape the static array serves, 16-bit input, plain store
    # call 2: right after it, fewer columns and blocks (stale partials of call 1 stay in the array), fp32, accumulate
    cases = [(300000, MAX_COLS, MAX_COLS, torch.bfloat16, 0), (1000, 3, 64, torch.float32, 1),
             (77, 500, 500, torch.float16, 0)]
    for rows, cols, ld, dtype, accumulate in cases:
        m = torch.randn(rows, ld, generator=g).to(dev).to(dtype)
        init = torch.randn(cols, generator=g).to(dev) if accumulate else None
        code = {torch.float32: lib.F32, torch.bfloat16: lib.BF16, torch.float16: lib.F16}[dtype]
        out_ws, out_st, (a0, f0, a1, f1) = _both(lib, m, rows, cols, ld, code, accumulate, init)
        ref = m.float()[:, :cols].double().sum(0)
        if init is not None:
            ref = ref + init.double()
        bound = TOL * float(m.float().abs().double().sum(0).max())
        assert float((out_st.double() - ref).abs().max()) < bound
        assert float((out_st - out_ws).abs().max()) < bound
        assert torch.equal(out_st, out_ws)  # same kernels, same partial layout, same summation order: same bits
        assert a1 == a0 and f1 == f0, "odvae_colsum_static changed the device's free memory"


def test_colsum_static_rejects_more_columns_than_the_array_holds():
    from odvae_b200 import lib
    lib.require_device()
    dev = torch.device("cuda", 0)
    cols = MAX_COLS + 64
    m = torch.zeros(128, cols, device=dev)
    out = torch.full((cols,), 7.0, device=dev)
    with pytest.raises(RuntimeError, match="ODVAE_COLSUM_STATIC_MAX_COLS"):
        lib.call("colsum_static", m, 128, cols, cols, lib.F32, out, 0)
    torch.cuda.synchronize()
    assert float(out.min()) == 7.0 and float(out.max()) == 7.0  # nothing was launched
