"""ngp_compact_alive_static: the allocation-free variant of ngp_compact_alive (scratch in a file-scope device array
inside the library; include/ngp_b200.h).  No counterpart in the reference.
Named test_zz_* so that pytest collects it last.

Checked through the C ABI against the existing entry point, bit-exact (integer ids and counts), on two calls in a row
with different inputs (the shared scratch keeps the previous call's contents), with and without the device-side count,
plus the size limit and that torch's allocator is not touched by the call."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _alive(n, keep, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    ids = torch.randperm(n, device="cuda", generator=g).int()
    dead = torch.rand(n, device="cuda", generator=g) >= keep
    return torch.where(dead, torch.full_like(ids, -1), ids).contiguous()


def _reference(abi, src, n, n_dev):
    out = torch.full((src.numel(),), -7, dtype=torch.int32, device="cuda")
    n_out = torch.zeros(1, dtype=torch.int32, device="cuda")
    ws = torch.zeros(max(1 << 16, abi.lib.ngp_compact_alive_workspace_bytes(n)), dtype=torch.uint8, device="cuda")
    abi.check(abi.lib.ngp_compact_alive(src.data_ptr(), n, abi.ptr(n_dev), out.data_ptr(), n_out.data_ptr(), ws.data_ptr(), ws.numel(), abi.stream()))
    return out, n_out


def test_compact_alive_static_matches_compact_alive_twice():
    import ngp_abi as abi
    torch.cuda.set_device(0)
    cases = [(300001, 0.6, None), (1920 * 1080, 0.05, 1234567), (1000, 1.0, None), (5, 0.0, None), (abi.COMPACT_STATIC_MAX_N, 0.5, None)]
    for rep in range(2):  # the whole list twice: every call after the first finds the scratch as another call left it
        for i, (n, keep, lim) in enumerate(cases):
            src = _alive(n, keep, seed=10 * rep + i)
            n_dev = torch.tensor([lim], dtype=torch.int32, device="cuda") if lim is not None else None
            ref_out, ref_n = _reference(abi, src, n, n_dev)
            out, out2 = (torch.full((n,), -7, dtype=torch.int32, device="cuda") for _ in range(2))
            n_out, n_out2 = (torch.zeros(1, dtype=torch.int32, device="cuda") for _ in range(2))
            torch.cuda.synchronize()
            before = (torch.cuda.memory_allocated(), torch.cuda.memory_reserved())
            abi.check(abi.lib.ngp_compact_alive_static(src.data_ptr(), n, abi.ptr(n_dev), out.data_ptr(), n_out.data_ptr(), abi.stream()))
            # second call straight after, same stream, same arguments: must give the same answer again
            abi.check(abi.lib.ngp_compact_alive_static(src.data_ptr(), n, abi.ptr(n_dev), out2.data_ptr(), n_out2.data_ptr(), abi.stream()))
            torch.cuda.synchronize()
            assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == before  # torch's allocator was not touched by the calls
            k = int(ref_n.item())
            head = src[: (lim if lim is not None else n)]
            assert k == int((head >= 0).sum().item())
            assert int(n_out.item()) == k and int(n_out2.item()) == k
            assert torch.equal(out, ref_out) and torch.equal(out2, ref_out)  # bit-exact, untouched tail included
            assert torch.equal(out[:k], head[head >= 0])  # stable order


def test_compact_alive_static_n_out_may_alias_n_dev_and_limit_is_enforced():
    import ngp_abi as abi
    torch.cuda.set_device(0)
    n = 50000
    src = _alive(n, 0.3, seed=99)
    n_dev = torch.tensor([40000], dtype=torch.int32, device="cuda")
    ref_out, ref_n = _reference(abi, src, n, n_dev.clone())
    out = torch.full((n,), -7, dtype=torch.int32, device="cuda")
    abi.check(abi.lib.ngp_compact_alive_static(src.data_ptr(), n, n_dev.data_ptr(), out.data_ptr(), n_dev.data_ptr(), abi.stream()))
    assert int(n_dev.item()) == int(ref_n.item()) and torch.equal(out, ref_out)
    # beyond the fixed scratch: a clear error, nothing launched
    big = abi.COMPACT_STATIC_MAX_N + 1
    with pytest.raises(abi.NgpError, match="NGP_COMPACT_STATIC_MAX_N"):
        abi.check(abi.lib.ngp_compact_alive_static(src.data_ptr(), big, None, out.data_ptr(), n_dev.data_ptr(), abi.stream()))
    with pytest.raises(abi.NgpError, match="aliasing"):
        abi.check(abi.lib.ngp_compact_alive_static(src.data_ptr(), n, None, src.data_ptr(), n_dev.data_ptr(), abi.stream()))
    # the library still works afterwards
    abi.check(abi.lib.ngp_compact_alive_static(src.data_ptr(), n, None, out.data_ptr(), n_dev.data_ptr(), abi.stream()))
    assert int(n_dev.item()) == int((src >= 0).sum().item())
