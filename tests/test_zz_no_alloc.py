"""gsr_compact_indices_noalloc (built-in static device scratch, allocates nothing) against gsr_compact_indices (caller
scratch) through the C-ABI: same indices, same count, bit for bit (integer work), on two calls in a row that share the
static array.  The entry point has no counterpart in the reference
(include/gsrast.h).  Collected last on purpose (file name)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _with_scratch(lib_mod, mask, add):
    P = mask.numel()
    nb = ctypes.c_size_t()
    lib_mod.check(lib_mod.lib.gsr_compact_scratch_bytes(P, ctypes.byref(nb)))
    scratch = torch.empty(nb.value, dtype=torch.uint8, device="cuda")
    idx = torch.full((max(P, 1),), -7, dtype=torch.int32, device="cuda")
    count = torch.full((1,), -1, dtype=torch.int64, device="cuda")
    lib_mod.check(lib_mod.lib.gsr_compact_indices(P, mask.data_ptr(), idx.data_ptr(), add, count.data_ptr(),
                                                  scratch.data_ptr(), _stream()))
    return idx, count


def _mask(P, density, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.rand(P, generator=g, device="cuda") < density).view(torch.uint8).contiguous()


def test_noalloc_matches_existing_entry_point_twice():
    from diff_gaussian_rasterization import _lib
    lib = _lib.lib
    # two calls in a row; the second is SMALLER and sparser, so the static array still holds the first call's block
    # offsets beyond the words the second call rewrites
    calls = [(1_000_003, 0.37, 11, 0), (70_001, 0.05, 12, 1_000_003)]
    masks = [_mask(P, d, s) for (P, d, s, _) in calls]
    want = [_with_scratch(_lib, m, add) for m, (_, _, _, add) in zip(masks, calls)]
    outs = [(torch.full((m.numel(),), -7, dtype=torch.int32, device="cuda"),
             torch.full((1,), -1, dtype=torch.int64, device="cuda")) for m in masks]
    torch.cuda.synchronize()
    free_before = None
    for k, (m, (P, _, _, add), (idx, count)) in enumerate(zip(masks, calls, outs)):
        if k == 1:  # every buffer exists and the module is loaded: the second call must not move free memory
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
        _lib.check(lib.gsr_compact_indices_noalloc(P, m.data_ptr(), idx.data_ptr(), add, count.data_ptr(), _stream()))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before, "the no-alloc entry point changed the device's free memory"
    for m, (P, _, _, add), (idx, count), (widx, wcount) in zip(masks, calls, outs, want):
        n = int(count.item())
        assert n == int(wcount.item()) == int(m.sum().item())
        assert torch.equal(idx, widx[:P]), "indices differ from gsr_compact_indices (bit-exact bar)"
        assert torch.equal(idx[:n].long() - add, m.bool().nonzero().squeeze(1))  # the order of tensor[mask]
        assert bool((idx[n:] == -7).all())  # nothing written past the count


def test_noalloc_edges_and_limit():
    from diff_gaussian_rasterization import _lib
    lib = _lib.lib
    count = torch.full((1,), -1, dtype=torch.int64, device="cuda")
    # P = 0: count 0, no pointers needed
    _lib.check(lib.gsr_compact_indices_noalloc(0, None, None, 0, count.data_ptr(), _stream()))
    assert int(count.item()) == 0
    # all set / none set / ragged last block
    for P, fill in ((4097, 1), (4096, 0), (1, 1)):
        m = torch.full((P,), fill, dtype=torch.uint8, device="cuda")
        idx = torch.full((P,), -7, dtype=torch.int32, device="cuda")
        _lib.check(lib.gsr_compact_indices_noalloc(P, m.data_ptr(), idx.data_ptr(), 5, count.data_ptr(), _stream()))
        n = int(count.item())
        assert n == P * fill
        assert torch.equal(idx[:n], torch.arange(n, dtype=torch.int32, device="cuda") + 5)
    # beyond the built-in scratch: a clear error, nothing launched
    before = _lib.launch_count()
    rc = lib.gsr_compact_indices_noalloc((1 << 24) + 1, count.data_ptr(), count.data_ptr(), 0, count.data_ptr(),
                                         _stream())
    assert rc != 0 and b"exceeds the built-in scratch" in lib.gsr_last_error()
    assert _lib.launch_count() == before
    # the largest accepted size runs (16 Mi mask bytes)
    P = 1 << 24
    m = _mask(P, 0.5, 3)
    idx = torch.empty(P, dtype=torch.int32, device="cuda")
    _lib.check(lib.gsr_compact_indices_noalloc(P, m.data_ptr(), idx.data_ptr(), 0, count.data_ptr(), _stream()))
    n = int(count.item())
    assert n == int(m.sum().item())
    assert torch.equal(idx[:n].long(), m.bool().nonzero().squeeze(1))
