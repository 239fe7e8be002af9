"""Fused-SGD semantics vs torch.optim.SGD, flat-arena invariants, DDP bucket planner."""
import torch

import distributed_pytorch_b200 as dpb
from distributed_pytorch_b200.parallel.ddp import expected_ready_order, plan_buckets
from distributed_pytorch_b200.utils.arena import FlatArena, adopt_module


def test_sgd_matches_torch_on_cpu():
    torch.manual_seed(0)
    a = [torch.nn.Parameter(torch.randn(5, 3)), torch.nn.Parameter(torch.randn(7))]
    b = [torch.nn.Parameter(p.detach().clone()) for p in a]
    oa = torch.optim.SGD(a, lr=0.1, momentum=0.9, weight_decay=1e-4)
    ob = dpb.SGD(b, lr=0.1, momentum=0.9, weight_decay=1e-4)
    for _ in range(4):
        for p, q in zip(a, b):
            g = torch.randn_like(p)
            p.grad, q.grad = g.clone(), g.clone()
        oa.step()
        ob.step()
    for p, q in zip(a, b):
        assert torch.allclose(p, q, rtol=1e-6, atol=1e-7)
    sa, sb = oa.state_dict(), ob.state_dict()
    assert sa["state"].keys() == sb["state"].keys()
    assert torch.allclose(sa["state"][0]["momentum_buffer"], sb["state"][0]["momentum_buffer"])


def test_arena_views_keep_schema_and_values():
    torch.manual_seed(1)
    m = dpb.VGG11()
    before = {k: v.clone() for k, v in m.state_dict().items()}
    arena = adopt_module(m)
    assert arena.total % (4 * 8) == 0 and all(o % 64 == 0 for o in arena.offsets)
    assert arena.names[0] == "layers.0.weight" and arena.names[-1] == "fc1.bias"
    for k, v in m.state_dict().items():
        assert torch.equal(v, before[k]) and v.shape == before[k].shape
    # parameters alias the flat buffer
    p = dict(m.named_parameters())["layers.4.weight"]
    p.data.fill_(2.0)
    i = arena.names.index("layers.4.weight")
    assert float(arena.param_buf[arena.offsets[i]]) == 2.0
    assert p.stride() == (576, 1, 192, 64)  # KRSC
    # gradient slot hands out fresh aliases of the gradient buffer
    g1, g2 = p._dpb_grad_slot(), p._dpb_grad_slot()
    assert g1 is not g2 and g1.data_ptr() == g2.data_ptr() == arena.grad_buf.data_ptr() + 4 * arena.offsets[i]
    assert arena.ensure_grads_in_place() == 34 and arena.grads_in_place()


def test_ready_order_and_bucket_layout_match_reference_ddp():
    m = dpb.VGG11()
    order = expected_ready_order(m)
    assert order[:6] == [33, 32, 30, 31, 28, 29] and order[-4:] == [2, 3, 0, 1] and sorted(order) == list(range(34))
    params = list(m.parameters())
    sizes = [params[i].numel() * 4 for i in order]
    ids = plan_buckets(sizes)
    totals = [sum(s for s, b in zip(sizes, ids) if b == k) for k in range(max(ids) + 1)]
    assert totals == [9461800, 27157504, 305152]  # rebuilt_bucket_sizes measured on torch DDP
    members = [[order[j] for j, b in enumerate(ids) if b == k] for k in range(3)]
    assert members[0] == [33, 32, 30, 31, 28]
    assert members[2] == [9, 6, 7, 4, 5, 2, 3, 0, 1]
    assert plan_buckets([100], 1 << 20, 25 << 20) == [0]


def test_reverse_bucket_plan_gives_contiguous_arena_ranges():
    """all_reduce mode with overlap keeps its arena in named_parameters() order; its buckets (cut back to front) must
    each cover one contiguous run of parameters, for every cap, or the reducer's arena ranges would interleave."""
    import distributed_pytorch_b200 as dpb
    from distributed_pytorch_b200.parallel.ddp import plan_reverse_buckets
    for name in ("VGG11", "VGG19"):
        sizes = [p.numel() * 4 for p in dpb.VGG(name).parameters()]
        for cap_mb in (1, 4, 8, 10, 25, 1000):
            ids = plan_reverse_buckets(sizes, 1 << 20, cap_mb << 20)
            assert len(ids) == len(sizes) and ids[-1] == 0          # the last parameter is in the first-ready bucket
            assert all(a >= b for a, b in zip(ids, ids[1:]))          # monotone => every bucket is one contiguous run
            assert sorted(set(ids)) == list(range(max(ids) + 1))
        assert max(plan_reverse_buckets(sizes, 1 << 20, 1000 << 20)) <= 1
