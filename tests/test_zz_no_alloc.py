"""The scan entry point that brings its own scratch, for callers that cannot allocate.

Some users of this library only ever hold the input and output buffers another component
gave them - a plugin in a CUDA context whose memory was divided up at start-up, a serving
loop that must not pay cudaMalloc's device-wide stall, a foreign-language binding with no
device allocator. They cannot give kmls_exclusive_scan_i32 the scratch it asks for, so
kmls_exclusive_scan_i32_noalloc uses a fixed array inside the library instead. This test
drives it through the C ABI and holds it to the regular entry point's output, exactly
(integers), on two calls in a row: the second call finds the first call's leftovers in the
shared array and must not be disturbed by them. The file name sorts last on purpose, so a
run that stops at the first failure has already run everything else.
"""
import numpy as np
import pytest
import torch

from kmls_b200 import _lib

pytestmark = pytest.mark.gpu
MAX_N = 4_194_304  # KMLS_NOALLOC_SCAN_MAX_N in include/kmls_b200.h


def _reference_scan(x: torch.Tensor) -> torch.Tensor:
    n = x.numel()
    out = torch.empty(n + 1, dtype=torch.int64, device=x.device)
    sb = int(_lib.value("kmls_scan_scratch_bytes", n))
    scratch = torch.empty(sb, dtype=torch.uint8, device=x.device)
    _lib.call("kmls_exclusive_scan_i32", x, n, out, scratch, sb, torch.cuda.current_stream().cuda_stream)
    return out


def _noalloc_scan(x: torch.Tensor) -> torch.Tensor:
    n = x.numel()
    out = torch.full((n + 1,), -1, dtype=torch.int64, device=x.device)
    _lib.call("kmls_exclusive_scan_i32_noalloc", x, n, out, torch.cuda.current_stream().cuda_stream)
    return out


def test_two_calls_in_a_row_match_the_regular_entry_point():
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(123)
    # first call: the largest size allowed; second call: a smaller, ragged size with other
    # data, so the array still holds tile states of the first call beyond what it resets
    for n in (MAX_N, 1_000_003):
        x = torch.from_numpy(rng.integers(0, 2000, n).astype(np.int32)).to(dev)
        want = _reference_scan(x)
        got = _noalloc_scan(x)
        torch.cuda.synchronize()
        assert torch.equal(got, want)
        assert int(got[-1].item()) == int(x.sum(dtype=torch.int64).item())


def test_empty_input_and_oversize_are_handled():
    dev = torch.device("cuda", 0)
    x = torch.zeros(1, dtype=torch.int32, device=dev)
    out = torch.full((1,), -1, dtype=torch.int64, device=dev)
    _lib.call("kmls_exclusive_scan_i32_noalloc", x, 0, out, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert out.tolist() == [0]
    with pytest.raises(_lib.KmlsError, match="exceeds the built-in scratch limit"):
        _lib.call("kmls_exclusive_scan_i32_noalloc", x, MAX_N + 1, out,
                  torch.cuda.current_stream().cuda_stream)
