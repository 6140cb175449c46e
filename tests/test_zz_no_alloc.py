"""asscpp_uncovered_ids_noalloc (include/asscpp_gpu.h): the allocation-free variant of the uncovered-id extraction,
scratch in a static device array.  No counterpart in the reference; checked here against the existing entry point
asscpp_uncovered_voxels on the same covered sets, bit for bit, on calls in a row that share the scratch array.

Collected last on purpose (the file name).
"""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_VOXELS = 1 << 23  # ASSCPP_NOALLOC_MAX_VOXELS
ERANGE = -4


@pytest.fixture(scope="module")
def A(lib_built):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import asscpp_b200
    return asscpp_b200


def call_noalloc(L, bits_dev, n_voxels, ids_dev, count_dev, stream):
    err = C.create_string_buffer(256)
    rc = L.asscpp_uncovered_ids_noalloc(C.c_void_p(bits_dev.data_ptr()), n_voxels, C.c_void_p(ids_dev.data_ptr()),
                                        ids_dev.numel(), C.c_void_p(count_dev.data_ptr()), C.c_void_p(stream), err, 256)
    assert rc == 0, err.value.decode()
    return rc


def test_noalloc_equals_existing_entry_point_twice_in_a_row(A, etihad, space_2p5):
    import torch
    ctx = A.CoverageContext(etihad[0], A.default_params())
    L = ctx.L
    n_cov = ctx.n_cov_voxels
    poses = np.ascontiguousarray(np.concatenate([space_2p5["cam0"][::7], space_2p5["cam1"][::7]]))
    s = ctx.new_set()
    stream = torch.cuda.current_stream().cuda_stream
    ids_dev = torch.full((n_cov,), -1, dtype=torch.int32, device="cuda")
    count_dev = torch.zeros(1, dtype=torch.int64, device="cuda")
    seen = []
    # many covered, then few covered (more ids than the call before: stale offsets in the scratch must not leak),
    # then nothing covered, then the whole sample again
    for k, cut in enumerate((len(poses), len(poses) // 2, 0, len(poses))):
        s.clear()
        if cut:
            s.or_views(poses[:cut], 5)
        want, _ = s.uncovered()                                    # asscpp_uncovered_voxels_ex, the existing entry point
        _, vox = s.export()
        bits_dev = torch.from_numpy(vox.view(np.int32).copy()).cuda()
        ids_dev.fill_(-1)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        call_noalloc(L, bits_dev, n_cov, ids_dev, count_dev, stream)
        torch.cuda.synchronize()
        if k > 0:  # (the first call may page the module in)
            assert torch.cuda.mem_get_info()[0] == free_before
        n = int(count_dev.cpu()[0])
        got = ids_dev.cpu().numpy()
        assert n == len(want)
        assert np.array_equal(got[:n], want)
        assert np.all(got[n:] == -1)
        seen.append(n)
    assert seen[0] == seen[3] and seen[0] <= seen[1] <= seen[2] == n_cov and seen[0] < n_cov
    s.free()
    ctx.close()


def test_noalloc_large_bitset_capped_and_refused_beyond_maximum(A):
    import torch
    from asscpp_b200 import _abi
    L = _abi.load()
    rng = np.random.default_rng(20261105)
    stream = torch.cuda.current_stream().cuda_stream
    count_dev = torch.zeros(1, dtype=torch.int64, device="cuda")
    for n_vox, cap in ((MAX_VOXELS, MAX_VOXELS), (1000003, 5000), (31, 31), (0, 0)):
        words = rng.integers(0, 2 ** 32, size=(n_vox + 31) // 32 + 1, dtype=np.uint64).astype(np.uint32)
        bits = np.unpackbits(words.view(np.uint8), bitorder="little")[:n_vox]
        want = np.nonzero(bits == 0)[0].astype(np.int32)
        bits_dev = torch.from_numpy(words.view(np.int32).copy()).cuda()
        ids_dev = torch.full((max(cap, 1),), -1, dtype=torch.int32, device="cuda")
        err = C.create_string_buffer(256)
        rc = L.asscpp_uncovered_ids_noalloc(C.c_void_p(bits_dev.data_ptr()), n_vox, C.c_void_p(ids_dev.data_ptr()), cap,
                                            C.c_void_p(count_dev.data_ptr()), C.c_void_p(stream), err, 256)
        assert rc == 0, err.value.decode()
        assert int(count_dev.cpu()[0]) == len(want)
        take = min(cap, len(want))
        assert np.array_equal(ids_dev.cpu().numpy()[:take], want[:take])
    # beyond the stated maximum: refused before anything is launched, with a clear text
    err = C.create_string_buffer(256)
    rc = L.asscpp_uncovered_ids_noalloc(C.c_void_p(bits_dev.data_ptr()), MAX_VOXELS + 1, C.c_void_p(ids_dev.data_ptr()), 1,
                                        C.c_void_p(count_dev.data_ptr()), C.c_void_p(stream), err, 256)
    assert rc == ERANGE
    assert b"ASSCPP_NOALLOC_MAX_VOXELS" in err.value and str(MAX_VOXELS).encode() in err.value
