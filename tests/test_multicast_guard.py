"""The condition under which the library uses TMA multicast loads (the A-multicast plan, cfg_cta_group = 4).

TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory
hung the GPU.  The library therefore takes that plan only after `tlb200_is_own_device_memory` has PROVEN that the
whole extent of A is one cudaMalloc allocation of the current device.  These tests exercise the predicate and the
plan decision only: none of the negative cases launches a kernel.
"""
import ctypes

import pytest
import torch

from tilelang_mesh_b200 import _abi
from tilelang_mesh_b200 import gemm as G

MC_SHAPE = (1000, 3072, 5120)  # a shape whose automatic plan is the A-multicast one when A is provably local


def test_plan_query_without_pointers_never_reports_multicast_unasked():
    # CPU: the pointer-free query cannot know where A lives
    for (m, n, k) in [MC_SHAPE, (4096, 4096, 4096), (8192, 8192, 8192), (2048, 2048, 2048)]:
        for dt in (torch.bfloat16, torch.float8_e4m3fn):
            assert G.plan(m, n, k, dt, torch.bfloat16)["multicast"] == 1
    assert G.plan(*MC_SHAPE, cta_group=4, block_n=256)["multicast"] == 2  # asked for by name


def _default_allocator_is_cudamalloc():
    import os
    conf = os.environ.get("PYTORCH_CUDA_ALLOC_CONF", "") + os.environ.get("PYTORCH_ALLOC_CONF", "")
    return "expandable_segments:True" not in conf and "cudaMallocAsync" not in conf


@pytest.mark.gpu
def test_predicate_on_this_gpus_memory():
    G._ensure_init(torch.cuda.current_device())
    if not _default_allocator_is_cudamalloc():
        pytest.skip("torch is not allocating with cudaMalloc in this process")
    t = torch.empty(1 << 20, dtype=torch.uint8, device="cuda")
    assert G.is_own_device_memory(t.data_ptr(), t.numel())
    assert G.is_own_device_memory(t.data_ptr() + 4096, 1000)           # inside an allocation
    assert not G.is_own_device_memory(t.data_ptr(), 1 << 40)           # runs past the end of the allocation
    assert not G.is_own_device_memory(t.data_ptr(), 0)
    assert not G.is_own_device_memory(0, 16)
    # two separate cudaMalloc allocations, even if adjacent in the address space, are not one extent
    big = [torch.empty(64 << 20, dtype=torch.uint8, device="cuda") for _ in range(2)]  # > 20 MiB: one cudaMalloc each
    lo, hi = sorted(x.data_ptr() for x in big)
    assert not G.is_own_device_memory(lo, hi - lo + 4096)


@pytest.mark.gpu
def test_predicate_rejects_everything_that_is_not_plain_local_device_memory():
    G._ensure_init(torch.cuda.current_device())
    pinned = torch.empty(1 << 16, dtype=torch.uint8).pin_memory()
    assert not G.is_own_device_memory(pinned.data_ptr(), pinned.numel())
    pageable = torch.empty(1 << 16, dtype=torch.uint8)
    assert not G.is_own_device_memory(pageable.data_ptr(), pageable.numel())
    buf = ctypes.create_string_buffer(4096)  # an address CUDA has never heard of
    assert not G.is_own_device_memory(ctypes.addressof(buf), 4096)
    assert not G.is_own_device_memory(0x7000_0000_0000, 4096)

    cudart = pytest.importorskip("cuda.bindings.runtime")
    driver = pytest.importorskip("cuda.bindings.driver")
    # managed memory
    err, p = cudart.cudaMallocManaged(1 << 16, cudart.cudaMemAttachGlobal)
    if err == cudart.cudaError_t.cudaSuccess:
        try:
            assert not G.is_own_device_memory(int(p), 1 << 16)
        finally:
            cudart.cudaFree(p)
    # stream-ordered pool memory
    err, p = cudart.cudaMallocAsync(1 << 16, 0)
    if err == cudart.cudaError_t.cudaSuccess:
        try:
            cudart.cudaStreamSynchronize(0)
            assert not G.is_own_device_memory(int(p), 1 << 16)
        finally:
            cudart.cudaFreeAsync(p, 0)
            cudart.cudaStreamSynchronize(0)
    # a virtual-memory mapping (cuMemCreate + cuMemMap), even of THIS GPU's physical memory, is never trusted:
    # such ranges are how symmetric memory and multi-GPU fabrics put peers' memory behind one address range
    dev = torch.cuda.current_device()
    prop = driver.CUmemAllocationProp()
    prop.type = driver.CUmemAllocationType.CU_MEM_ALLOCATION_TYPE_PINNED
    prop.location.type = driver.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
    prop.location.id = dev
    err, gran = driver.cuMemGetAllocationGranularity(prop, driver.CUmemAllocationGranularity_flags.CU_MEM_ALLOC_GRANULARITY_MINIMUM)
    if err == driver.CUresult.CUDA_SUCCESS:
        size = int(gran)
        err, h = driver.cuMemCreate(size, prop, 0)
        assert err == driver.CUresult.CUDA_SUCCESS
        err, va = driver.cuMemAddressReserve(size, 0, 0, 0)
        assert err == driver.CUresult.CUDA_SUCCESS
        try:
            assert driver.cuMemMap(va, size, 0, h, 0)[0] == driver.CUresult.CUDA_SUCCESS
            acc = driver.CUmemAccessDesc()
            acc.location.type = driver.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
            acc.location.id = dev
            acc.flags = driver.CUmemAccess_flags.CU_MEM_ACCESS_FLAGS_PROT_READWRITE
            assert driver.cuMemSetAccess(va, size, [acc], 1)[0] == driver.CUresult.CUDA_SUCCESS
            assert not G.is_own_device_memory(int(va), size)
            driver.cuMemUnmap(va, size)
        finally:
            driver.cuMemAddressFree(va, size)
            driver.cuMemRelease(h)


@pytest.mark.gpu
def test_plan_follows_the_predicate_and_explicit_request_fails_closed():
    """Plan decision only (tlb200_gemm_plan_query_call launches nothing); the one launch below uses local memory."""
    G._ensure_init(torch.cuda.current_device())
    if not _default_allocator_is_cudamalloc():
        pytest.skip("torch is not allocating with cudaMalloc in this process")
    m, n, k = MC_SHAPE
    a = torch.randint(-2, 3, (m, k), device="cuda").float().bfloat16()
    b = torch.randint(-2, 3, (n, k), device="cuda").float().bfloat16()
    c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    assert G.plan_for(a, b, c, cta_group=4, block_n=256)["multicast"] == 2
    auto = G.plan_for(a, b, c)
    # A that the device can read but that is not its own memory: pinned host memory (mapped into the device's
    # address space).  Same shape, same everything: the plan must not be the multicast one, with or without room
    # for split-K in the workspace (the fall-back to an unsplit schedule re-resolves the plan).
    a_host = torch.empty(m, k, dtype=torch.bfloat16).pin_memory()
    for ws_bytes in (None, 256):
        p = G.plan_for(a_host, b, c, workspace_bytes=ws_bytes)
        assert p["multicast"] == 1, p
    with pytest.raises(RuntimeError, match="cfg_cta_group=4"):
        G.plan_for(a_host, b, c, cta_group=4, block_n=256)
    # the launching entry point makes the same decision before anything is launched
    lib = _abi.load()
    n0 = lib.tlb200_launch_count()
    desc = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, k, k, n, cta_group=4, block_n=256)
    rc = lib.tlb200_gemm(ctypes.byref(desc), a_host.data_ptr(), b.data_ptr(), c.data_ptr(), None)
    assert rc != 0 and "cfg_cta_group=4" in _abi.last_error()
    assert lib.tlb200_launch_count() == n0
    # a view whose extent leaves its allocation is not provably local either (pitch says rows are further apart than
    # they are: the tensor map would reach past the end of the allocation)
    desc_far = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, 1 << 30, k, n, cta_group=4, block_n=256)
    rc = lib.tlb200_gemm(ctypes.byref(desc_far), a.data_ptr(), b.data_ptr(), c.data_ptr(), None)
    assert rc != 0 and lib.tlb200_launch_count() == n0
    # local memory: the multicast plan runs and is right
    want = (a.float() @ b.float().t()).bfloat16()
    got = G.gemm(a, b, cta_group=4, block_n=256)
    assert torch.equal(got, want)
    if auto["multicast"] == 2:
        assert torch.equal(G.gemm(a, b), want)


@pytest.mark.gpu
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
def test_predicate_rejects_peer_memory():
    """Predicate only: nothing is launched on either GPU."""
    G._ensure_init(0)
    with torch.cuda.device(1):
        peer = torch.empty(1 << 20, dtype=torch.uint8, device="cuda:1")
    with torch.cuda.device(0):
        assert not G.is_own_device_memory(peer.data_ptr(), peer.numel())
        if torch.cuda.can_device_access_peer(0, 1):
            # enable peer access the way torch does for cross-device copies, then ask again
            torch.empty(16, device="cuda:0").copy_(peer[:16].view(torch.uint8).to(torch.float32)[:16], non_blocking=False)
            assert not G.is_own_device_memory(peer.data_ptr(), peer.numel())
        m, n, k = MC_SHAPE
        a_peer = torch.empty(m, k, dtype=torch.bfloat16, device="cuda:1")
        b = torch.empty(n, k, dtype=torch.bfloat16, device="cuda:0")
        c = torch.empty(m, n, dtype=torch.bfloat16, device="cuda:0")
        lib = _abi.load()
        out = _abi.GemmPlan()
        desc = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, k, k, n)
        assert lib.tlb200_gemm_plan_query_call(ctypes.byref(desc), a_peer.data_ptr(), b.data_ptr(), c.data_ptr(), None, 0,
                                               ctypes.byref(out)) == 0
        assert out.multicast == 1
        desc4 = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, k, k, n, cta_group=4, block_n=256)
        assert lib.tlb200_gemm_plan_query_call(ctypes.byref(desc4), a_peer.data_ptr(), b.data_ptr(), c.data_ptr(), None, 0,
                                               ctypes.byref(out)) != 0
