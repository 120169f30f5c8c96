"""gsb_radix_sort_pairs_u32_noalloc (scratch in a static device array inside the library; no counterpart in the
reference) against gsb_radix_sort_pairs_u32, through the C ABI.

Collected last on purpose (the file name sorts behind every other test file).
Integer work: the bar is bit-exact, as for every sort / intersection test of this repo."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _sort(entry, keys, vals, end_bit, n_dev=None):
    """One call of either entry point on copies of (keys, vals); -> (sorted keys, permuted vals)."""
    from gsplat.b200_impl._lib import lib, ptr, stream_of

    k, v = keys.clone(), vals.clone()
    k_alt, v_alt = torch.empty_like(k), torch.empty_like(v)
    in_alt = ctypes.c_int(0)
    n = k.numel()
    if entry == "noalloc":
        lib.gsb_radix_sort_pairs_u32_noalloc(n, n_dev, end_bit, ptr(k), ptr(k_alt), ptr(v), ptr(v_alt),
                                             ctypes.addressof(in_alt), stream_of(k))
    else:
        temp_bytes = int(lib.gsb_radix_temp_bytes(n, end_bit))
        temp = torch.empty((max(temp_bytes, 1),), dtype=torch.uint8, device=k.device)
        lib.gsb_radix_sort_pairs_u32(n, n_dev, end_bit, ptr(k), ptr(k_alt), ptr(v), ptr(v_alt), ptr(temp), temp_bytes,
                                     ctypes.addressof(in_alt), stream_of(k))
    torch.cuda.synchronize()
    return (k_alt, v_alt) if in_alt.value else (k, v)


def _case(n, end_bit, seed):
    g = torch.Generator().manual_seed(seed)
    hi = (1 << end_bit) if end_bit < 31 else (1 << 31) - 1
    keys = torch.randint(0, hi, (n,), generator=g, dtype=torch.int64).to(torch.int32)
    keys[::5] = keys[0]  # ties: the sort is stable, the values tell
    return keys.to(DEV), torch.arange(n, dtype=torch.int32, device=DEV)


def test_noalloc_sort_matches_the_temp_buffer_sort_on_two_calls_in_a_row():
    # two calls back to back on the shared scratch: a large one (many tiles: look-back states of every pass used),
    # then a smaller one with other bit widths, which must not see anything the first left behind
    for n, end_bit, seed in ((1_500_000, 32, 1), (70_001, 13, 2)):
        keys, vals = _case(n, end_bit, seed)
        k_ref, v_ref = _sort("temp", keys, vals, end_bit)
        k_new, v_new = _sort("noalloc", keys, vals, end_bit)
        assert torch.equal(k_new, k_ref) and torch.equal(v_new, v_ref)
        # and against torch (stable sort of the unsigned bit patterns)
        order = torch.sort(keys.to(torch.int64) & 0xFFFFFFFF, stable=True).indices
        assert torch.equal(v_new.long(), order)


def test_noalloc_sort_device_side_count_and_limits():
    from gsplat.b200_impl._lib import lib, ptr, stream_of

    keys, vals = _case(300_000, 20, 3)
    n_dev = torch.tensor([123_457], dtype=torch.int64, device=DEV)  # only the leading items are sorted
    k_ref, v_ref = _sort("temp", keys, vals, 20, ptr(n_dev))
    k_new, v_new = _sort("noalloc", keys, vals, 20, ptr(n_dev))
    m = int(n_dev.item())
    assert torch.equal(k_new[:m], k_ref[:m]) and torch.equal(v_new[:m], v_ref[:m])
    # empty input is fine; one item beyond the built-in scratch is refused with a clear message, nothing launched
    in_alt = ctypes.c_int(0)
    lib.gsb_radix_sort_pairs_u32_noalloc(0, None, 32, None, None, None, None, ctypes.addressof(in_alt), stream_of(keys))
    too_many = (1 << 21) + 1
    with pytest.raises(RuntimeError, match="GSB_RADIX_NOALLOC_MAX_ITEMS"):
        lib.gsb_radix_sort_pairs_u32_noalloc(too_many, None, 32, ptr(keys), ptr(keys), ptr(vals), ptr(vals),
                                             ctypes.addressof(in_alt), stream_of(keys))
