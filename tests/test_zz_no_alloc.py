"""GPU: the allocation-free sort entry point (trase_b200_sort_pairs_static).

Some hosts of this library cannot allocate device memory when they sort — they
gave the whole GPU to their own arena, run under a hard memory cap, or must keep
cudaMalloc/cudaFree (which can stall every stream) out of their frame loop.  For
them the library carries its own fixed histogram workspace.  This test drives
that entry point through the C ABI and requires its output to be identical to
the regular entry point's (integer work: bit-exact), on two calls in a row so
that whatever the first call left in the shared workspace cannot leak into the
second.  The file sorts last on purpose (pytest collects it after the other
tests), and it must pass both on its own and after them.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

STATIC_MAX = 1 << 22


def _pairs(n, cap, seed, end_bit):
    rng = np.random.default_rng(seed)
    keys = (rng.integers(0, 5440, cap, dtype=np.uint64) << np.uint64(32)) | rng.integers(0, 500, cap, dtype=np.uint64)
    keys &= np.uint64((1 << end_bit) - 1)        # only the low end_bit bits take part in the sort
    vals = rng.permutation(cap).astype(np.uint32)
    return keys, vals


def _run(native_lib, which, keys, vals, n, cap, end_bit, dev):
    k_in = torch.from_numpy(keys.view(np.int64).copy()).to(dev)
    v_in = torch.from_numpy(vals.view(np.int32).copy()).to(dev)
    k_out, v_out = torch.zeros_like(k_in), torch.zeros_like(v_in)
    cnt = torch.tensor([n], dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    if which == "regular":
        sb = native_lib.trase_b200_sort_scratch_bytes(cap)
        scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
        rc = native_lib.trase_b200_sort_pairs(k_in.data_ptr(), v_in.data_ptr(), k_out.data_ptr(), v_out.data_ptr(),
                                              cnt.data_ptr(), cap, end_bit, scratch.data_ptr(), sb, stream)
    else:
        # every buffer above was allocated before this call; the call itself allocates nothing
        rc = native_lib.trase_b200_sort_pairs_static(k_in.data_ptr(), v_in.data_ptr(), k_out.data_ptr(),
                                                     v_out.data_ptr(), cnt.data_ptr(), cap, end_bit, stream)
    assert rc == 0, native_lib.trase_b200_last_error()
    torch.cuda.synchronize(dev)
    return k_out[:n].cpu().numpy(), v_out[:n].cpu().numpy()


def test_static_sort_equals_regular_sort_twice_in_a_row(native_lib):
    dev = torch.device("cuda:0")
    # two different problems back to back: a large one, then a smaller one with
    # another digit distribution and an odd pass count, reusing the same workspace
    for n, cap, end_bit, seed in [(1_500_000, 1_500_017, 45, 1), (123_457, 200_000, 39, 2)]:
        keys, vals = _pairs(n, cap, seed, end_bit)
        k_ref, v_ref = _run(native_lib, "regular", keys, vals, n, cap, end_bit, dev)
        k_new, v_new = _run(native_lib, "static", keys, vals, n, cap, end_bit, dev)
        assert np.array_equal(k_new, k_ref)
        assert np.array_equal(v_new, v_ref)
        order = np.argsort(keys[:n], kind="stable")          # and both are the stable sort
        assert np.array_equal(k_new.view(np.uint64), keys[:n][order])
        assert np.array_equal(v_new.view(np.uint32), vals[:n][order])


def _check(native_lib, n, cap, end_bit, seed, dev):
    keys, vals = _pairs(n, cap, seed, end_bit)
    k_ref, v_ref = _run(native_lib, "regular", keys, vals, n, cap, end_bit, dev)
    k_new, v_new = _run(native_lib, "static", keys, vals, n, cap, end_bit, dev)
    assert np.array_equal(k_new, k_ref)
    assert np.array_equal(v_new, v_ref)
    order = np.argsort(keys[:n], kind="stable")
    assert np.array_equal(k_new.view(np.uint64), keys[:n][order])
    assert np.array_equal(v_new.view(np.uint32), vals[:n][order])


def test_static_sort_even_pass_count_and_full_workspace(native_lib):
    """end_bit 34 takes four 9-bit passes, so the result comes back through the copy that follows an even pass
    count; the second case fills the built-in workspace to its documented limit."""
    dev = torch.device("cuda:0")
    _check(native_lib, 1_200_000, 1_200_000, 34, 3, dev)
    _check(native_lib, STATIC_MAX, STATIC_MAX, 45, 4, dev)


def test_static_sort_big_small_big_reuses_stale_workspace(native_lib):
    """The workspace keeps the previous call's histograms, and sorts up to 2^20 pairs use a different tile size
    (and therefore a different histogram layout) than larger ones: alternate them."""
    dev = torch.device("cuda:0")
    for n, cap, end_bit, seed in [(2_000_000, 2_000_000, 45, 5), (50_001, 60_000, 36, 6), (1_100_000, 1_100_000, 45, 7)]:
        _check(native_lib, n, cap, end_bit, seed, dev)


def test_static_sort_rejects_null_buffers_and_clears_old_errors(native_lib):
    dev = torch.device("cuda:0")
    rc = native_lib.trase_b200_sort_pairs_static(None, None, None, None, None, 1000, 45, None)
    assert rc != 0
    assert "null buffer" in native_lib.trase_b200_last_error().decode()
    _check(native_lib, 1000, 1000, 40, 8, dev)                # a success after a failure ...
    assert native_lib.trase_b200_last_error().decode() == ""  # ... leaves no stale message behind


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs a second GPU in this process")
def test_static_sort_on_a_second_gpu_of_the_same_process(native_lib):
    """The workspace address and the kernels' shared-memory opt-in are per GPU: the same process sorting on
    cuda:1 after cuda:0 must work (a sort above 2^20 pairs takes the kernel that needs the opt-in)."""
    _check(native_lib, 1_300_000, 1_300_000, 45, 9, torch.device("cuda:0"))
    with torch.cuda.device(1):
        _check(native_lib, 1_300_000, 1_300_000, 45, 10, torch.device("cuda:1"))


def test_static_sort_refuses_sizes_beyond_its_workspace(native_lib):
    rc = native_lib.trase_b200_sort_pairs_static(None, None, None, None, None, STATIC_MAX + 1, 45, None)
    assert rc != 0
    msg = native_lib.trase_b200_last_error().decode()
    assert "at most %d" % STATIC_MAX in msg
