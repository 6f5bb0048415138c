"""The depth sort on its own (csrc/onesweep.cuh and csrc/depth_sort.cuh through gsr_sort_pairs) against
torch.sort(stable=True): stable ascending order over all 32 key bits, values = original indices.  Semantics of the
reference's cub::DeviceRadixSort::SortPairs call (DGR/cuda_rasterizer/rasterizer_impl.cu:300-308)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _sort(keys, variant):
    from diff_gaussian_rasterization import _lib
    n = keys.numel()
    nb = ctypes.c_size_t()
    _lib.check(_lib.lib.gsr_sort_scratch_bytes(n, ctypes.byref(nb)))
    scratch = torch.empty(nb.value, dtype=torch.uint8, device="cuda")
    k0 = keys.clone()
    k1, v0, v1 = (torch.empty_like(keys) for _ in range(3))
    _lib.check(_lib.lib.gsr_sort_pairs(n, k0.data_ptr(), k1.data_ptr(), v0.data_ptr(), v1.data_ptr(),
                                       scratch.data_ptr(), variant, torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return k0, v0


@pytest.mark.parametrize("variant", [1, 0])
@pytest.mark.parametrize("n,kind", [(1, "rand"), (31, "rand"), (4096, "rand"), (4097, "dup"), (33730, "depth"),
                                    (100_003, "dup"), (1_000_000, "depth"), (2_500_001, "rand"), (50_000, "same")])
def test_sort_pairs_is_a_stable_32_bit_sort(n, kind, variant):
    g = torch.Generator(device="cuda").manual_seed(n)
    if kind == "rand":      # all 32 bits used, incl. the sign bit
        keys = torch.randint(-2**31, 2**31 - 1, (n,), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    elif kind == "dup":     # few distinct keys: stability decides the order of the values
        keys = torch.randint(0, 7, (n,), generator=g, device="cuda", dtype=torch.int64).to(torch.int32) * 0x01010101
    elif kind == "same":    # one digit value in every pass
        keys = torch.full((n,), 0x40A00000, device="cuda", dtype=torch.int32)
    else:                   # depth-like: positive floats in [0.2, 10) with 14 % "culled" 0xffffffff markers
        z = 0.2 + 9.8 * torch.rand(n, generator=g, device="cuda")
        keys = z.view(torch.int32).clone()
        keys[torch.rand(n, generator=g, device="cuda") < 0.14] = -1
    want_k, want_v = torch.sort(keys.to(torch.int64) & 0xFFFFFFFF, stable=True)
    k, v = _sort(keys, variant)
    assert torch.equal(k.to(torch.int64) & 0xFFFFFFFF, want_k)
    assert torch.equal(v.to(torch.int64), want_v)
