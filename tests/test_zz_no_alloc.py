"""sp_materialise_haplotype_static (no counterpart in the reference): the
no-allocation variant of haplotype materialisation, whose plan tables live in a file-scope device array of the library.
Called through the C ABI and compared bit for bit with the existing entry points (sp_plan_haplotypes + sp_materialise +
sp_copy_haplotype), on calls in a row that share the array. Collected last on purpose."""
import numpy as np
import pytest

import oracle_helpers as oh

pytestmark = pytest.mark.gpu


def _existing(ctx, h):
    lens, stride = ctx.plan_haplotypes(h, h, False)
    ctx.materialise()
    return ctx.copy_haplotype(h, stride)


def test_static_scratch_variant_matches_the_existing_entry_points(sp):
    import torch
    rng = np.random.default_rng(20261105)
    case = oh.random_edit_lists(rng, 3_000_000, 900, 4, max_len=4000)
    case["lists"][3] = np.zeros(0, np.int32)   # a haplotype without edits: the reference itself
    list_off = np.concatenate(([0], np.cumsum([x.size for x in case["lists"]]))).astype(np.int64)
    ids = np.concatenate(case["lists"]).astype(np.int32)
    with sp.Context(0, 42) as ctx:
        ctx.set_reference(case["ref"])
        ctx.set_variants(case["sv_type"], case["position"], case["length"], case["payload_off"], case["payload"])
        ctx.set_haplotype_lists(list_off, ids)
        want = [_existing(ctx, h).copy() for h in range(4)]
        assert len({w.size for w in want}) > 1 and want[3].tobytes() == case["ref"].tobytes()
        capacity = max(w.size for w in want) + 64
        out = torch.full((capacity,), 0x2A, dtype=torch.uint8, device="cuda")
        # the longest list first, then shorter ones: tables left by a call must not leak into the next
        order = sorted(range(4), key=lambda h: -case["lists"][h].size) + [0, 0]
        for h in order:
            d_ids = torch.from_numpy(case["lists"][h].copy()).to("cuda") if case["lists"][h].size else torch.zeros(1, dtype=torch.int32, device="cuda")
            out.fill_(0x2A)
            torch.cuda.synchronize()
            l0 = ctx.kernel_launches
            n = ctx.materialise_haplotype_static(d_ids.data_ptr(), int(case["lists"][h].size), out.data_ptr(), capacity)
            assert ctx.kernel_launches - l0 == 6   # init, plan, count, scan, fill, copy
            got = out.cpu().numpy()
            assert n == want[h].size, (h, n, want[h].size)
            assert got[:n].tobytes() == want[h].tobytes(), h
            assert (got[n:] == 0x2A).all(), "bytes past the haplotype were written"
        # the existing path is untouched by the calls in between
        assert _existing(ctx, 1).tobytes() == want[1].tobytes()


def test_static_scratch_variant_rejects_what_does_not_fit(sp):
    import torch
    rng = np.random.default_rng(7)
    case = oh.random_edit_lists(rng, 200_000, 50, 1, max_len=500)
    with sp.Context(0, 42) as ctx:
        ctx.set_reference(case["ref"])
        ctx.set_variants(case["sv_type"], case["position"], case["length"], case["payload_off"], case["payload"])
        out = torch.empty(1 << 20, dtype=torch.uint8, device="cuda")
        d_ids = torch.from_numpy(case["lists"][0].copy()).to("cuda")
        torch.cuda.synchronize()
        n_ids = int(case["lists"][0].size)
        with pytest.raises(sp.SynthpopError, match="SP_STATIC_MAX_EDITS = 16384"):
            ctx.materialise_haplotype_static(d_ids.data_ptr(), 16385, out.data_ptr(), out.numel())
        with pytest.raises(sp.SynthpopError, match="too small"):
            ctx.materialise_haplotype_static(d_ids.data_ptr(), n_ids, out.data_ptr(), 1000)
        with pytest.raises(sp.SynthpopError, match="16-byte aligned"):
            ctx.materialise_haplotype_static(d_ids.data_ptr(), n_ids, out.data_ptr() + 4, 1 << 19)
        # and still works afterwards
        n = ctx.materialise_haplotype_static(d_ids.data_ptr(), n_ids, out.data_ptr(), out.numel())
        ctx.set_haplotype_lists(np.array([0, n_ids], np.int64), case["lists"][0])
        assert out.cpu().numpy()[:n].tobytes() == _existing(ctx, 0).tobytes()
