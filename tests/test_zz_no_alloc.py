"""-m gpu, collected last: tutor_lmhead_argmax_static, the LM-head arg-max for callers that cannot allocate
device memory (their own pool owns the GPU heap, or they must stay off the driver's allocation lock).  It must
return exactly what tutor_lmhead_argmax returns - both run the same kernels, only the home of the scratch
differs - on two calls in a row: the built-in scratch array is shared and keeps the previous call's contents,
so the second call (other inputs, other batch size) would expose any dependence on stale data."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _case(seed, B, V, d, dtype):
    g = torch.Generator().manual_seed(seed)
    td = torch.float32 if dtype == 0 else torch.bfloat16
    x = torch.randn(B, d, generator=g).to(DEV).to(td)
    wte = (torch.randn(V, d, generator=g) * 0.02).to(DEV).to(td)
    seen = torch.randint(-2**31, 2**31 - 1, (B, (V + 31) // 32), generator=g, dtype=torch.int64).to(torch.int32).to(DEV)
    return x, wte, seen


@pytest.mark.parametrize("dtype", [0, 1], ids=["fp32", "bf16"])
def test_static_scratch_matches_allocating_entry_point_twice(pkg, lib, dtype):
    check = pkg._lib.check
    # second call: smaller batch and vocabulary than the first, so stale partials from call 1 sit right behind it
    for seed, B, V, d, penalty in ((1, 64 if dtype else 8, 50257, 768, 1.2), (2, 5, 3000, 128, 1.0)):
        x, wte, seen = _case(seed, B, V, d, dtype)
        ref = torch.empty(B, dtype=torch.int32, device=DEV)
        scratch = torch.empty(lib.tutor_lmhead_scratch_bytes(B), dtype=torch.uint8, device=DEV)
        check(lib.tutor_lmhead_argmax(dtype, x.data_ptr(), wte.data_ptr(), B, V, d, seen.data_ptr(), penalty,
                                      ref.data_ptr(), scratch.data_ptr(), None))
        out = torch.full((B,), -7, dtype=torch.int32, device=DEV)
        check(lib.tutor_lmhead_argmax_static(dtype, x.data_ptr(), wte.data_ptr(), B, V, d, seen.data_ptr(), penalty,
                                             out.data_ptr(), None))
        torch.cuda.synchronize()
        assert torch.equal(out, ref), (seed, B)
        assert int(out.min()) >= 0 and int(out.max()) < V


def test_static_scratch_rejects_oversize_batch(pkg, lib):
    x, wte, seen = _case(3, 129, 1000, 128, 1)
    out = torch.empty(129, dtype=torch.int32, device=DEV)
    rc = lib.tutor_lmhead_argmax_static(1, x.data_ptr(), wte.data_ptr(), 129, 1000, 128, seen.data_ptr(), 1.0,
                                        out.data_ptr(), None)
    assert rc == -4  # TUTOR_ERR_LIMIT
    msg = lib.tutor_last_error().decode()
    assert "129" in msg and "128" in msg
