"""rn_masks_expand_static (the entry point that allocates nothing: scratch in a __device__ array of librn_cuda) against
rn_masks_expand on the same masks, through the C ABI.  Integer work: the lists and counts must be identical.
Collected last on purpose (the file name sorts after every other test file)."""
import numpy as np
import pytest

from rayne_b200 import capi

pytestmark = pytest.mark.gpu


def _segments(torch, seed, sizes, density):
    """Random mask words per segment (device int32 tensors) and their rank bases: consecutive ranges with a gap.
    Also the list the bits stand for, built on the host."""
    rng = np.random.Generator(np.random.PCG64(seed))
    tensors, segs, want, base = [], [], [], 7
    for nw in sizes:
        bits = rng.random(nw * 32) < density
        words = np.packbits(bits, bitorder="little").view(np.uint32) if nw else np.zeros(0, np.uint32)
        t = torch.from_numpy(np.concatenate([words, np.zeros(1, np.uint32)]).view(np.int32)).cuda()
        tensors.append(t)
        segs.append((t.data_ptr(), nw, base))
        want.append(np.nonzero(bits)[0].astype(np.uint32) + np.uint32(base))
        base += nw * 32 + 96
    return tensors, segs, np.concatenate(want)


def _run(torch, fn, segs, capacity, start):
    out = torch.full((capacity + 8,), -1, dtype=torch.int32, device="cuda")
    cnt = torch.full((1,), start, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    fn(segs, out.data_ptr(), capacity, cnt.data_ptr())
    torch.cuda.synchronize()
    return out.cpu().numpy(), int(cnt.item())


def test_static_expand_matches_the_allocating_entry_point_twice_in_a_row():
    import torch
    # two different inputs back to back: the second call finds the first one's totals in the shared array (a longer call
    # first, so a shorter one would read stale entries if it read past its own), then the first input again
    cases = [
        dict(seed=1, sizes=[70001, 1, 0, 1023, 4096], density=0.3),
        dict(seed=2, sizes=[513, 2050], density=0.9),
        dict(seed=1, sizes=[70001, 1, 0, 1023, 4096], density=0.3),
        dict(seed=3, sizes=[300000], density=0.02),
    ]
    with capi.SceneHandle(max_nodes=16) as h:
        for case in cases:
            tensors, segs, want = _segments(torch, **case)
            for capacity, start in ((len(want) + 5, 0), (len(want) // 2, 0), (len(want) + 5, 3)):
                got_s, cnt_s = _run(torch, capi.masks_expand_static, segs, capacity, start)
                got_a, cnt_a = _run(torch, h.masks_expand, segs, capacity, start)
                assert cnt_s == cnt_a == start + len(want)
                assert np.array_equal(got_s, got_a), case
                k = max(min(capacity - start, len(want)), 0)
                assert np.array_equal(got_s[start:start + k].astype(np.uint32), want[:k])
                assert (got_s[:start] == -1).all() and (got_s[start + k:] == -1).all()


def test_static_expand_allocates_no_device_memory():
    import torch
    tensors, segs, want = _segments(torch, seed=5, sizes=[200000, 777], density=0.5)
    out = torch.full((len(want),), -1, dtype=torch.int32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    capi.masks_expand_static(segs, out.data_ptr(), len(want), cnt.data_ptr())      # first call: the address look-up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(2):
        cnt.zero_()
        capi.masks_expand_static(segs, out.data_ptr(), len(want), cnt.data_ptr(), device=torch.cuda.current_device())
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free0 == free1
    assert int(cnt.item()) == len(want)
    assert np.array_equal(out.cpu().numpy().astype(np.uint32), want)


def test_static_expand_rejects_what_exceeds_its_array():
    import torch
    w = torch.zeros(4, dtype=torch.int32, device="cuda")
    out = torch.zeros(4, dtype=torch.int32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    limit = capi.MASKS_EXPAND_STATIC_MAX_WORDS
    with pytest.raises(capi.RNError) as e:
        capi.masks_expand_static([(w.data_ptr(), limit + 1, 0)], out.data_ptr(), 4, cnt.data_ptr())   # rejected before any launch
    assert e.value.status == -5 and "RN_MASKS_EXPAND_STATIC_MAX_WORDS" in str(e.value)
    with pytest.raises(capi.RNError) as e:
        capi.masks_expand_static([(w.data_ptr(), limit // 2, 0), (w.data_ptr(), limit // 2 + 1, 0)], out.data_ptr(), 4, cnt.data_ptr())
    assert e.value.status == -5
    with pytest.raises(capi.RNError) as e:
        capi.masks_expand_static([(w.data_ptr(), 1, 0)] * 33, out.data_ptr(), 4, cnt.data_ptr())
    assert e.value.status == -1
    torch.cuda.synchronize()
    assert int(cnt.item()) == 0
    # exactly at the limit is accepted (all-zero words: nothing is written)
    big = torch.zeros(limit, dtype=torch.int32, device="cuda")
    capi.masks_expand_static([(big.data_ptr(), limit, 0)], out.data_ptr(), 4, cnt.data_ptr())
    torch.cuda.synchronize()
    assert int(cnt.item()) == 0
