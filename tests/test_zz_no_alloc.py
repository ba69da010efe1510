"""irf_depth_scan_static — the allocation-free depth scan — against the regular path (irf_finalize), through the C ABI.

Some hosts of this library cannot call cudaMalloc once they are running: their GPU memory is handed out in full at
start-up, or they cannot tolerate the device-wide synchronisation an allocation causes while other streams are busy.
For them the scan's scratch lives in a fixed array inside the library.  That array is shared by every call and keeps
its old contents, so this test makes two calls in a row, on different data, and requires both to match exactly the
depths the regular entry point produces from the same difference counters (integer work: the tolerance is zero).
The file name sorts last on purpose: a failure here cannot hide a failure anywhere else under `pytest -x`.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_ELEMS = 1 << 27


def _counted_core(tmp_path, name, seed, pairs):
    from irfinder_b200 import IRFinderCore, synth
    meta = synth.generate(os.path.join(str(tmp_path), name), genome="tiny", introns=150, pairs=pairs, readlen=100, seed=seed, chunks=3, collated=False)
    core = IRFinderCore(0)
    core.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
    data = np.fromfile(meta["sorted"], dtype=np.uint8)
    core.stage(data.ctypes.data, data.size)
    core.count()          # difference counters are final, depths not computed yet
    return core


def test_static_scratch_scan_matches_finalize_twice_in_a_row(built, tmp_path):
    import torch
    from irfinder_b200 import IRFError, depth_scan_static
    from irfinder_b200.distributed import _wrap
    dev = torch.device("cuda", 0)
    results = []
    for name, seed, pairs in (("first", 51, 9000), ("second", 52, 5000)):     # second call: other data, shorter array
        core = _counted_core(tmp_path, name, seed, pairs)
        ptr, n, _ = core.device_table("diff")
        assert 0 < n <= MAX_ELEMS
        diff = _wrap(ptr, n, "<i4", dev).clone()                              # the caller's own input buffer
        out = torch.full_like(diff, -12345)                                   # and its own output buffer
        depth_scan_static(diff.data_ptr(), out.data_ptr(), n)                 # new entry point, default stream
        torch.cuda.synchronize()
        core.finalize()                                                       # existing entry point scans the context's copy
        want = core.table("depth")
        got = out.cpu().numpy()
        assert got.shape == want.shape and np.array_equal(got, want), name
        assert want.max() > 0                                                 # the comparison is not vacuous
        inplace = diff.clone()
        depth_scan_static(inplace.data_ptr(), inplace.data_ptr(), n)          # same buffer in and out is allowed
        torch.cuda.synchronize()
        assert np.array_equal(inplace.cpu().numpy(), want), name + " (in place)"
        results.append(n)
        core.close()
    assert results[0] != results[1]
    # a request beyond the fixed scratch is refused with a message that names the limit, and nothing is launched
    dummy = torch.zeros(16, dtype=torch.int32, device=dev)
    with pytest.raises(IRFError) as ei:
        depth_scan_static(dummy.data_ptr(), dummy.data_ptr(), MAX_ELEMS + 1)
    assert ei.value.code == -6 and str(MAX_ELEMS) in str(ei.value)
    depth_scan_static(dummy.data_ptr(), dummy.data_ptr(), 0)                  # empty input is a no-op
