"""GPU: f5e_grn_static, the allocation-free variant of f5e_grn (include/f5e_b200.h; no counterpart in the reference - added
for callers that may not call cudaMalloc).  Its scratch is a file-scope __device__ array of the library, so the test checks,
through the C ABI, on two calls in a row: same bits as f5e_grn (same kernels), the torch formula of model/modules.py:225-234
within the tolerance of tests/test_kernels_gpu.py, no allocation by torch's allocator during the calls, and a clear error
beyond the built-in limit.  Named zz so pytest collects it last."""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_FLOATS = 1 << 20  # F5E_GRN_STATIC_MAX_FLOATS


def rel_l2(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-20)).item()


def _case(dev, seed, B, N, Cc):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(B, N, Cc, device=dev, generator=g).bfloat16()
    gamma, beta = torch.randn(Cc, device=dev, generator=g), torch.randn(Cc, device=dev, generator=g)
    return x, gamma, beta


def _torch_grn(x, gamma, beta):
    xf = x.float()
    Gx = xf.norm(p=2, dim=1, keepdim=True)
    Nx = Gx / (Gx.mean(dim=-1, keepdim=True) + 1e-6)
    return gamma * (xf * Nx) + beta + xf


def test_grn_static_matches_grn_twice(cuda_dev):
    from f5e_tts_b200 import _lib, ops
    from f5e_tts_b200.ops import check, ptr, stream_ptr
    lib = _lib.load()
    # two different shapes back to back: the second call finds the first one's partials in the shared array
    for seed, (B, N, Cc) in enumerate([(3, 90, 1024), (2, 333, 512)]):
        x, gamma, beta = _case(cuda_dev, 100 + seed, B, N, Cc)
        want = ops.grn(x, gamma, beta)                       # existing entry point, caller-provided scratch
        ref = _torch_grn(x, gamma, beta)
        assert lib.f5e_grn_static_scratch_floats(B, N, Cc) == B * ((N + 63) // 64) * Cc <= MAX_FLOATS
        out = torch.full_like(x, float("nan"))
        for call in range(2):
            torch.cuda.synchronize()
            n_alloc = torch.cuda.memory_stats()["allocation.all.allocated"]
            free0, _ = torch.cuda.mem_get_info()
            check(lib.f5e_grn_static(ptr(x), ptr(gamma), ptr(beta), ptr(out), B, N, Cc, stream_ptr()), "f5e_grn_static")
            torch.cuda.synchronize()
            assert torch.cuda.memory_stats()["allocation.all.allocated"] == n_alloc   # nothing from torch's allocator ...
            if call == 1:  # ... nor from the driver.  (The very first call of a process may see the driver materialise the
                # 4 MiB array itself: CUDA loads the module of csrc/grn_static.cu lazily, at its first use.)
                assert torch.cuda.mem_get_info()[0] == free0
            assert torch.equal(out, want), (seed, call)       # same kernels, fixed summation order: same bits
            assert rel_l2(out, ref) < 4e-3, (seed, call)      # the bar of tests/test_kernels_gpu.py
            out.fill_(float("nan"))


def test_grn_static_rejects_oversize(cuda_dev):
    from f5e_tts_b200 import _lib
    from f5e_tts_b200.ops import ptr, stream_ptr
    lib = _lib.load()
    B, N, Cc = 17, 64 * 64, 1024                             # 17 * 64 * 1024 floats > 2^20
    assert lib.f5e_grn_static_scratch_floats(B, N, Cc) > MAX_FLOATS
    x = torch.zeros(1, 8, Cc, device=cuda_dev, dtype=torch.bfloat16)   # never touched: the size check comes first
    gamma = torch.zeros(Cc, device=cuda_dev)
    rc = lib.f5e_grn_static(ptr(x), ptr(gamma), ptr(gamma), ptr(x), B, N, Cc, stream_ptr())
    assert rc == -1
    msg = lib.f5e_last_error().decode()
    assert "F5E_GRN_STATIC_MAX_FLOATS" in msg and "f5e_grn" in msg
    # the largest shape that fits still works (exactly the limit: 16 samples x 64 slabs x 1024 channels)
    B, N = 16, 64 * 64
    assert lib.f5e_grn_static_scratch_floats(B, N, Cc) == MAX_FLOATS
    x, gamma, beta = _case(cuda_dev, 7, B, N, Cc)
    out = torch.empty_like(x)
    assert lib.f5e_grn_static(ptr(x), ptr(gamma), ptr(beta), ptr(out), B, N, Cc, stream_ptr()) == 0
    from f5e_tts_b200 import ops
    assert torch.equal(out, ops.grn(x, gamma, beta))
