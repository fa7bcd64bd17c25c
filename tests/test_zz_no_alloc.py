"""GPU test of nnd_query_noalloc_cuda_f32 (include/nnd_b200.h): the no-allocation variant of nnd_query_cuda_f32 whose
scratch is a static device array inside libnnd_b200.so.  Called through the C ABI (ctypes) and compared BIT-EXACTLY
(the tolerance every forward test in this suite uses: idx and fp32 dist identical) with the existing entry point, on two
calls in a row that share the array.  Collected last on purpose."""
import pytest

from _util import bits_equal

pytestmark = pytest.mark.gpu

NND_ETOOBIG = 10004


def _clouds(torch, device, b, nq, nt, seed):
    g = torch.Generator().manual_seed(seed)
    q = torch.rand(b, nq, 3, generator=g)
    t = torch.rand(b, nt, 3, generator=g)
    # ties on purpose: a block of query points are exact copies of target points, and targets repeat
    t[:, nt // 2:nt // 2 + 64] = t[:, :64]
    q[:, :128] = t[:, nt // 2 - 64:nt // 2 + 64]
    return q.to(device).contiguous(), t.to(device).contiguous()


def _existing(torch, lib, q, t):
    b, nq, _ = q.shape
    nt = t.shape[1]
    dist = torch.full((b, nq), -1.0, device=q.device)
    idx = torch.full((b, nq), -1, dtype=torch.int32, device=q.device)
    wsb = lib.nnd_query_workspace_bytes(b, nq, nt)
    ws = torch.empty(max(wsb, 256), dtype=torch.uint8, device=q.device)
    rc = lib.nnd_query_cuda_f32(q.data_ptr(), t.data_ptr(), b, nq, nt, dist.data_ptr(), idx.data_ptr(), ws.data_ptr(),
                                wsb, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.nnd_error_string(rc)
    torch.cuda.synchronize()
    return dist, idx


def test_noalloc_query_matches_the_existing_entry_point_twice(cuda_lib, device):
    import torch

    lib = cuda_lib
    assert lib.nnd_query_noalloc_max_bytes() == 8 << 20
    # shapes of the two-pass path (the one with scratch): second call smaller in one cloud and larger in the other, so
    # it runs over whatever the first call left in the array
    shapes = [(4, 6000, 5000, 11), (3, 4500, 7001, 12)]
    stream = torch.cuda.Stream(device=device)
    for call, (b, nq, nt, seed) in enumerate(shapes):
        assert 0 < lib.nnd_query_workspace_bytes(b, nq, nt) <= lib.nnd_query_noalloc_max_bytes()
        q, t = _clouds(torch, device, b, nq, nt, seed)
        want_dist, want_idx = _existing(torch, lib, q, t)
        dist = torch.full((b, nq), -1.0, device=device)
        idx = torch.full((b, nq), -1, dtype=torch.int32, device=device)
        torch.cuda.synchronize()
        allocated = torch.cuda.memory_allocated(device)
        free_before = torch.cuda.mem_get_info(device)[0] if call else None  # the first call may load the module
        with torch.cuda.stream(stream):
            rc = lib.nnd_query_noalloc_cuda_f32(q.data_ptr(), t.data_ptr(), b, nq, nt, dist.data_ptr(), idx.data_ptr(),
                                                stream.cuda_stream)
        assert rc == 0, lib.nnd_error_string(rc)
        assert lib.nnd_last_launch_count() == 3  # prepare + scan + finalize, all on the static scratch
        stream.synchronize()
        assert torch.cuda.memory_allocated(device) == allocated
        if free_before is not None:
            assert torch.cuda.mem_get_info(device)[0] == free_before, "the no-allocation call changed free device memory"
        assert bits_equal(idx.cpu().numpy(), want_idx.cpu().numpy()), f"call {call}: idx differs"
        assert bits_equal(dist.cpu().numpy(), want_dist.cpu().numpy()), f"call {call}: dist differs"
        # the planted copies are found at distance 0
        assert float(dist[:, :128].abs().max()) == 0.0


def test_noalloc_query_against_the_oracle(cuda_lib, device):
    import torch

    import oracle

    b, nq, nt = 4, 5000, 4200  # two-pass path (a lone or double pair of this size takes the single launch)
    q, t = _clouds(torch, device, b, nq, nt, 13)
    dist = torch.empty((b, nq), device=device)
    idx = torch.empty((b, nq), dtype=torch.int32, device=device)
    rc = cuda_lib.nnd_query_noalloc_cuda_f32(q.data_ptr(), t.data_ptr(), b, nq, nt, dist.data_ptr(), idx.data_ptr(),
                                             torch.cuda.current_stream().cuda_stream)
    assert rc == 0, cuda_lib.nnd_error_string(rc)
    assert cuda_lib.nnd_last_launch_count() == 3
    torch.cuda.synchronize()
    want_dist, want_idx = oracle.query(q.cpu().numpy(), t.cpu().numpy())
    assert bits_equal(idx.cpu().numpy(), want_idx) and bits_equal(dist.cpu().numpy(), want_dist)


def test_noalloc_query_limits_and_small_shapes(cuda_lib, device):
    import torch

    lib = cuda_lib
    # exactly at the limit is accepted by the size check, one tile more is not; nothing is launched or written
    assert lib.nnd_query_workspace_bytes(1, 262144, 262144) == lib.nnd_query_noalloc_max_bytes()
    b, nq, nt = 1, 262144 + 512, 262144
    q = torch.zeros(b, nq, 3, device=device)
    t = torch.zeros(b, nt, 3, device=device)
    dist = torch.full((b, nq), -7.0, device=device)
    idx = torch.full((b, nq), -7, dtype=torch.int32, device=device)
    rc = lib.nnd_query_noalloc_cuda_f32(q.data_ptr(), t.data_ptr(), b, nq, nt, dist.data_ptr(), idx.data_ptr(), None)
    assert rc == NND_ETOOBIG and lib.nnd_last_launch_count() == 0
    assert b"NND_NOALLOC_SCRATCH_BYTES" in lib.nnd_error_string(rc)
    torch.cuda.synchronize()
    assert float(dist.max()) == -7.0 and int(idx.max()) == -7
    assert lib.nnd_query_noalloc_cuda_f32(None, None, 1, -1, 4, None, None, None) == 10001
    # shapes of the single-launch kernel need no scratch and go through as well
    q, t = _clouds(torch, device, 8, 700, 300, 14)
    want_dist, want_idx = _existing(torch, lib, q, t)
    dist = torch.empty((8, 700), device=device)
    idx = torch.empty((8, 700), dtype=torch.int32, device=device)
    rc = lib.nnd_query_noalloc_cuda_f32(q.data_ptr(), t.data_ptr(), 8, 700, 300, dist.data_ptr(), idx.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    assert bits_equal(idx.cpu().numpy(), want_idx.cpu().numpy()) and bits_equal(dist.cpu().numpy(), want_dist.cpu().numpy())
    # an empty target cloud: dist = 0, idx = 0 (my_lib.cpp:9-10), like the existing entry point
    dist.fill_(3.0)
    rc = lib.nnd_query_noalloc_cuda_f32(q.data_ptr(), t.data_ptr(), 8, 700, 0, dist.data_ptr(), idx.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    assert float(dist.abs().max()) == 0.0 and int(idx.abs().max()) == 0
