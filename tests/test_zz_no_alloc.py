"""GPU: pim_snappy_compress_device_noalloc (scratch in the library's built-in device array, no allocation in the call;
no counterpart in the reference) gives exactly the bytes and block offsets
of pim_snappy_compress_device -- on two calls in a row that share the array (a larger input first, so that the second
call sees the first one's leftovers), and rejects inputs the array cannot serve.  Collected last on purpose
(`pytest --collect-only`).  Passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest


def _compress_both(P, torch, data, bs):
    L = P.lib()
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    n = data.size
    nb = int(L.pim_snappy_num_blocks(n, bs))
    d_in = torch.from_numpy(data.copy()).to(dev)
    bound = int(L.pim_snappy_compress_bound(n, bs))
    # reference point: the existing entry point with caller-provided scratch
    scratch = torch.empty(int(L.pim_snappy_compress_scratch_bytes(n, bs)), dtype=torch.uint8, device=dev)
    out_a = torch.full((bound,), 0x11, dtype=torch.uint8, device=dev)
    off_a = torch.zeros(nb + 1, dtype=torch.int64, device=dev)
    P.compress_device(d_in.data_ptr(), n, bs, scratch.data_ptr(), scratch.numel(), out_a.data_ptr(), off_a.data_ptr(), st)
    # the new entry point: outputs allocated BEFORE the call, free device memory compared across it
    out_b = torch.full((bound,), 0x22, dtype=torch.uint8, device=dev)
    off_b = torch.zeros(nb + 1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(dev)[0]
    P.compress_device_noalloc(d_in.data_ptr(), n, bs, out_b.data_ptr(), off_b.data_ptr(), st)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info(dev)[0]
    total = int(off_a[-1].item())
    return out_a[:total].cpu().numpy(), off_a.cpu().numpy(), out_b[:total].cpu().numpy(), off_b.cpu().numpy(), free0, free1


@pytest.mark.gpu
def test_noalloc_entry_point_matches_the_existing_one(P, oracle):
    import torch
    assert P.set_devices([0]) == 1
    calls = [(P.synth_host(P.SYNTH_MIX, 0x2A110C, (6 << 20) + 4321), 32768),   # first call: the larger input
             (P.synth_host(P.SYNTH_TEXT, 0x2A110D, (1 << 20) + 77), 4096),      # second call, other block size, leftovers in the array
             (P.synth_host(P.SYNTH_BIN, 0x2A110E, 300001), 65536)]
    frees = []
    for data, bs in calls:
        a, off_a, b, off_b, free0, free1 = _compress_both(P, torch, data, bs)
        assert np.array_equal(off_a, off_b), bs     # bit-exact: integer / byte work
        assert np.array_equal(a, b), bs
        want = oracle.compress(data, bs)              # and both are the oracle's stream body
        assert np.array_equal(b, want[want.size - b.size:]), bs
        frees.append((free0, free1))
    for free0, free1 in frees[1:]:                    # (the first call may load lazily; later ones must not take device memory)
        assert free1 >= free0, (free0, free1)


@pytest.mark.gpu
def test_noalloc_entry_point_rejects_what_the_array_cannot_serve(P):
    import torch
    L = P.lib()
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    n, bs = 64 << 20, 32768
    assert int(L.pim_snappy_compress_scratch_bytes(n, bs)) > P.api.NOALLOC_SCRATCH_BYTES
    d_in = torch.zeros(16, dtype=torch.uint8, device=dev)  # never touched: the size check comes first
    d_out = torch.full((16,), 0x33, dtype=torch.uint8, device=dev)
    d_off = torch.full((2,), -1, dtype=torch.int64, device=dev)
    with pytest.raises(P.SnappyError) as e:
        P.compress_device_noalloc(d_in.data_ptr(), n, bs, d_out.data_ptr(), d_off.data_ptr(), st)
    assert e.value.status == P.SNAPPY_BUFFER_TOO_SMALL
    with pytest.raises(P.SnappyError):
        P.compress_device_noalloc(d_in.data_ptr(), 16, 0, d_out.data_ptr(), d_off.data_ptr(), st)
    torch.cuda.synchronize()
    assert bool((d_out == 0x33).all().item()) and bool((d_off == -1).all().item())
    # the largest 32 KiB-block input the array serves is accepted
    n_ok = 16 << 20
    assert int(L.pim_snappy_compress_scratch_bytes(n_ok, bs)) <= P.api.NOALLOC_SCRATCH_BYTES
