"""dfn_compact_indices_static: stream compaction whose scratch is a fixed array inside the library (include/deepfields_b200.h).

Why the entry point exists: dfn_compact_indices wants a workspace from its caller, and not every caller can produce one - an
engine that reserved the whole GPU for its own arena makes a later cudaMalloc fail, a frame loop cannot afford the device-wide
stall an allocation brings, a plug-in is handed input / output buffers and may not create device memory.  This file checks,
through the C ABI, that the workspace-free variant returns exactly what dfn_compact_indices returns (index work: bit-exact),
twice in a row on the same GPU - the second call finds the scratch full of the first call's counts and ballots and must not be
disturbed by them - and that a list beyond DFN_COMPACT_STATIC_MAX_N is refused with a message instead of overrunning the array.

The file name sorts after every other test file on purpose: `pytest -x` stops at the first failure, so a failure here cannot
hide another test.  The test does not depend on anything the earlier tests did (it passes alone in a fresh process, where its
first call performs the per-GPU address look-up, and after the whole suite).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    assert torch.cuda.is_available()
    import deepfields_nerf_b200 as d

    return d._lib


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _with_workspace(lib, flags, idx_in):
    """the existing entry point: the caller (here: torch) allocates the workspace"""
    n = flags.numel()
    out = torch.full((max(n, 1),), -7, device="cuda", dtype=torch.int32)
    count = torch.full((1,), -1, device="cuda", dtype=torch.int64)
    ws = torch.empty(max(int(lib.load().dfn_compact_indices_workspace_bytes(n)), 16), device="cuda", dtype=torch.uint8)
    lib.call("dfn_compact_indices", flags.data_ptr(), idx_in.data_ptr() if idx_in is not None else None, n, out.data_ptr(),
             count.data_ptr(), ws.data_ptr(), _stream())
    return out, count


def _without_workspace(lib, flags, idx_in):
    """the new entry point: inputs and outputs only"""
    n = flags.numel()
    out = torch.full((max(n, 1),), -7, device="cuda", dtype=torch.int32)
    count = torch.full((1,), -1, device="cuda", dtype=torch.int64)
    lib.call("dfn_compact_indices_static", flags.data_ptr(), idx_in.data_ptr() if idx_in is not None else None, n, out.data_ptr(),
             count.data_ptr(), _stream())
    return out, count


def test_static_scratch_compaction_matches_workspace_compaction_twice(lib):
    g = torch.Generator(device="cuda").manual_seed(20261106)
    max_n = int(lib.load().dfn_compact_indices_static_max_n())
    assert max_n == 1 << 22
    # two calls in a row; the second is SHORTER and differently filled than the first, so every chunk count / ballot word the
    # first call left behind the second call's range is stale, and the ones inside it must be overwritten, not accumulated into
    cases = [
        (1920 * 1080, 0.37, True),    # the rays of one frame with an incoming index list (config 5's active list)
        (1000 * 1024 + 333, 0.81, False),   # ragged tail, positions themselves
    ]
    for n, keep, with_list in cases:
        flags = (torch.rand(n, device="cuda", generator=g) < keep).to(torch.uint8)
        idx_in = torch.randperm(n, device="cuda", generator=g).to(torch.int32) if with_list else None
        ref_out, ref_count = _with_workspace(lib, flags, idx_in)
        out, count = _without_workspace(lib, flags, idx_in)
        torch.cuda.synchronize()
        k = int(ref_count.item())
        assert int(count.item()) == k == int(flags.sum().item())
        assert torch.equal(out, ref_out)            # bit-exact, untouched tail (-7) included
        src = idx_in if with_list else torch.arange(n, device="cuda", dtype=torch.int32)
        assert torch.equal(out[:k], src[flags.bool()])
    # the limit itself is accepted ...
    flags = (torch.rand(max_n, device="cuda", generator=g) < 0.5).to(torch.uint8)
    ref_out, ref_count = _with_workspace(lib, flags, None)
    out, count = _without_workspace(lib, flags, None)
    assert int(count.item()) == int(ref_count.item()) and torch.equal(out, ref_out)
    # ... an empty list gives a zero count and launches nothing ...
    out, count = _without_workspace(lib, torch.empty(0, device="cuda", dtype=torch.uint8), None)
    assert int(count.item()) == 0
    # ... and one flag more than the limit is refused with a message that names the limit
    flags = torch.ones(max_n + 1, device="cuda", dtype=torch.uint8)
    with pytest.raises(RuntimeError, match="exceed the built-in scratch"):
        _without_workspace(lib, flags, None)
    torch.cuda.synchronize()
