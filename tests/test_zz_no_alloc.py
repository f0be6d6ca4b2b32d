"""artis_gpu_pack_escaped_noalloc (include/artis_gpu.h): the escaped-packet compaction with its scratch in a fixed
file-scope device array, for callers that cannot reach cudaMalloc. Checked bit for bit against the existing entry point
artis_gpu_packets_download_escaped on the same packets, on two calls in a row with different inputs (every call shares
the one scratch array, which keeps the previous call's contents). Collected last (test_zz_*).

The test owns the device buffers (torch tensors); the entry point under test allocates nothing."""
import numpy as np
import pytest

from artis_b200 import abi
from artis_b200.transport import ArtisGpuError, Transport

pytestmark = pytest.mark.gpu


def _packets(n, seed, frac_escaped):
    rng = np.random.default_rng(seed)
    raw = rng.integers(0, 256, size=n * abi.PACKET_DTYPE.itemsize, dtype=np.uint8)  # every byte of every record is checked
    pk = raw.view(abi.PACKET_DTYPE).copy()
    types = np.array([abi.TYPE_RPKT, abi.TYPE_KPKT, abi.TYPE_GAMMA, abi.TYPE_PELLET, abi.TYPE_MA], dtype=np.int32)
    pk["type"] = types[rng.integers(0, types.size, size=n)]
    pk["type"][rng.random(n) < frac_escaped] = abi.TYPE_ESCAPE
    pk["number"] = np.arange(n, dtype=np.int32)
    # the resident ensemble of the existing entry point keeps the two enums and the macro-atom element / ion in 16 bits
    # each (common.cuh PktHot): give them values of their own ranges, everything else stays random bytes
    pk["last_cross"] = rng.integers(101, 108, size=n)
    pk["ma_element"] = rng.integers(0, 30, size=n)
    pk["ma_ion"] = rng.integers(0, 10, size=n)
    return pk


def _reference_selection(T, pk):
    """the existing entry point: resident ensemble -> host array of the escaped packets"""
    T.packets_upload(pk)
    return T.download_escaped(only_new=False)


def test_pack_escaped_noalloc_matches_download_escaped_twice(gpu_lib):
    import torch
    T = Transport(gpu_lib)
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    import goldenlib
    fx = goldenlib.load_fixture("classic_nogrey_ts12.npz")
    T.init(fx["static"], device=0)  # the existing entry point works on the library's resident ensemble
    # two calls in a row: the second has fewer packets and another escape pattern, so that stale counts of the first
    # call are still lying in the scratch array beyond the chunks the second one writes
    cases = [(300007, 11, 0.31), (70001, 12, 0.02), (257, 13, 1.0), (1, 14, 0.0)]
    itemsize = abi.PACKET_DTYPE.itemsize
    for n, seed, frac in cases:
        pk = _packets(n, seed, frac)
        want = _reference_selection(T, pk)
        d_in = torch.from_numpy(pk.view(np.uint8).reshape(-1)).to(dev)
        d_out = torch.full((n * itemsize,), 0xEE, dtype=torch.uint8, device=dev)
        d_n = torch.full((1,), -7, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        T.pack_escaped_noalloc(d_in.data_ptr(), n, d_out.data_ptr(), n, d_n.data_ptr())
        torch.cuda.synchronize()
        got_n = int(d_n.cpu()[0])
        assert got_n == want.size == int((pk["type"] == abi.TYPE_ESCAPE).sum())
        got = d_out.cpu().numpy()[: got_n * itemsize].view(abi.PACKET_DTYPE)
        assert got.tobytes() == want.tobytes()  # bit-exact, padding bytes included
        # nothing beyond the packed records was touched
        assert (d_out.cpu().numpy()[got_n * itemsize:] == 0xEE).all()
    # a full buffer: only `capacity` records are written, the count is still the true one
    pk = _packets(5000, 21, 0.5)
    n_esc = int((pk["type"] == abi.TYPE_ESCAPE).sum())
    cap = n_esc // 3
    d_in = torch.from_numpy(pk.view(np.uint8).reshape(-1)).to(dev)
    d_out = torch.full((n_esc * itemsize,), 0xEE, dtype=torch.uint8, device=dev)
    d_n = torch.zeros(1, dtype=torch.int32, device=dev)
    T.pack_escaped_noalloc(d_in.data_ptr(), pk.size, d_out.data_ptr(), cap, d_n.data_ptr())
    torch.cuda.synchronize()
    assert int(d_n.cpu()[0]) == n_esc
    out = d_out.cpu().numpy()
    # whole records, padding included (a boolean index on the structured array itself would drop the padding bytes)
    want = pk.view(np.uint8).reshape(-1, itemsize)[pk["type"] == abi.TYPE_ESCAPE][:cap]
    assert out[: cap * itemsize].tobytes() == want.tobytes()
    assert (out[cap * itemsize:] == 0xEE).all()
    # beyond the fixed scratch: rejected with a clear error, nothing launched
    with pytest.raises(ArtisGpuError, match="exceed the fixed scratch"):
        T.pack_escaped_noalloc(d_in.data_ptr(), (1 << 22) + 1, d_out.data_ptr(), 1, d_n.data_ptr())
    T.finalize()
