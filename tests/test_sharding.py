"""CPU tests of the (view, scale) sharding plan and of the gradient / statistics exchange (gloo, world_size 2).
The rasterizer itself is not involved: these cover the host-side logic of ms-gs_b200/msgs_sharding.py."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import msgs_sharding as sh


def test_units_and_whole_view_assignment():
    units = sh.make_units(8, [1, 2, 4, 8, 16, 32], 1237, 822)
    assert len(units) == 48
    assert (units[1].W, units[1].H) == (618, 411) and (units[5].W, units[5].H) == (38, 25)   # int(W / s), int(H / s)
    for world in (1, 2, 4, 8):
        plan = sh.assign_units(units, world, P=3_000_000)
        assert sorted(i for r in plan for i in r) == list(range(48))            # every unit exactly once
        for r in plan:                                                           # whole views, identical unit mix
            assert len(r) == 48 // world
            assert {units[i].scale for i in r} == {1, 2, 4, 8, 16, 32}
            assert len({units[i].view for i in r}) == 8 // world


def test_lpt_assignment_when_views_do_not_divide():
    units = sh.make_units(3, [1, 2, 4, 8], 1237, 822)
    plan = sh.assign_units(units, 2, P=1_200_000)
    assert sorted(i for r in plan for i in r) == list(range(12))
    loads = [sum(sh.unit_cost(1_200_000, units[i]) for i in r) for r in plan]
    biggest = max(sh.unit_cost(1_200_000, u) for u in units)
    assert abs(loads[0] - loads[1]) <= biggest                                   # LPT bound
    # ragged: more ranks than units, and no units at all
    plan = sh.assign_units(units[:1], 4, P=10)
    assert sorted(map(len, plan)) == [0, 0, 0, 1]
    assert sh.assign_units([], 2, P=10) == [[], []]
    with pytest.raises(ValueError):
        sh.assign_units(units, 0, P=10)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_unit_grads(P, unit_index):
    g = torch.Generator().manual_seed(1000 + unit_index)
    return {name: torch.randn((P, n), generator=g) for name, n in sh.GRAD_FIELDS}


def _worker(rank, world, port, P, n_units, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        units = sh.make_units(n_units // 2, [1, 2], 64, 48)
        mine = sh.assign_units(units, world, P)[rank]
        ex = sh.GradientExchange(P, "cpu")
        st = sh.DensificationStats(P, levels=2, device="cpu")
        ex.zero()
        for i in mine:
            ex.add(_fake_unit_grads(P, i))
            g = torch.Generator().manual_seed(5000 + i)
            radii = torch.randint(0, 5, (P,), generator=g, dtype=torch.int32)
            px = torch.rand((P,), generator=g) * (torch.rand((P,), generator=g) > 0.3)
            d2 = torch.randn((P, 3), generator=g)
            st.update(0 if units[i].scale == 1 else 1, radii, px, d2)
        ex.reduce(async_op=True)
        red = ex.wait()
        st.reduce()
        if rank == 0:
            torch.save({"grads": {k: v.clone() for k, v in red.items()}, "stats": {
                "grad_accum": st.grad_accum, "denom": st.denom, "max_radii2D": st.max_radii2D, "px_min": st.px_min,
                "px_max": st.px_max, "seen": st.seen}, "mine": mine}, os.path.join(out_dir, "rank0.pt"))
        # every rank must hold the same reduced values
        chk = red["sh"].sum().reshape(1).clone()
        both = [torch.zeros(1) for _ in range(world)]
        dist.all_gather(both, chk)
        assert all(torch.equal(both[0], b) for b in both)
    finally:
        dist.destroy_process_group()


def test_gradient_and_stats_exchange_world_size_2(tmp_path):
    P, n_units, world = 257, 8, 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, P, n_units, str(tmp_path)), nprocs=world, join=True)
    got = torch.load(os.path.join(str(tmp_path), "rank0.pt"))
    # serial ground truth: one process doing every unit
    units = sh.make_units(n_units // 2, [1, 2], 64, 48)
    ex = sh.GradientExchange(P, "cpu")
    st = sh.DensificationStats(P, levels=2, device="cpu")
    for i in range(len(units)):
        ex.add(_fake_unit_grads(P, i))
        g = torch.Generator().manual_seed(5000 + i)
        radii = torch.randint(0, 5, (P,), generator=g, dtype=torch.int32)
        px = torch.rand((P,), generator=g) * (torch.rand((P,), generator=g) > 0.3)
        d2 = torch.randn((P, 3), generator=g)
        st.update(0 if units[i].scale == 1 else 1, radii, px, d2)
    for name, _ in sh.GRAD_FIELDS:
        assert torch.allclose(got["grads"][name], ex.views[name], rtol=1e-5, atol=1e-5), name
    assert torch.allclose(got["stats"]["grad_accum"], st.grad_accum, rtol=1e-5, atol=1e-6)
    assert torch.equal(got["stats"]["denom"], st.denom)
    assert torch.equal(got["stats"]["max_radii2D"], st.max_radii2D)
    assert torch.equal(got["stats"]["px_min"], st.px_min)
    assert torch.equal(got["stats"]["px_max"], st.px_max)
    assert torch.equal(got["stats"]["seen"], st.seen)
    assert len(got["mine"]) == n_units // world


def test_exchange_is_identity_without_process_group():
    ex = sh.GradientExchange(5, "cpu")
    g = _fake_unit_grads(5, 0)
    ex.add(g)
    ex.add(g)
    out = ex.reduce()
    assert torch.allclose(out["sh"].view(5, 48), 2 * g["sh"])
    assert ex.bytes_per_step() == 5 * 59 * 4
    with pytest.raises(ValueError):
        ex.add({k: v[:3] for k, v in g.items()})


def test_whole_view_assignment_yields_view_groups():
    """C4 (8 views x 6 scales): every rank's units are whole views with the scales consecutive, so that each view forms
    ONE view group (one preprocess launch + one depth sort + one preprocess backward per view, msgs_forward_group_begin)
    under the units API's view_ids."""
    import os
    import sys
    import torch
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(ROOT, "ms-gs_b200"))
    import diff_gaussian_rasterization as dgr
    import msgs_sharding
    units = msgs_sharding.make_units(8, [1, 2, 4, 8, 16, 32], 1237, 822)
    for world in (1, 2, 4, 8):
        plan = msgs_sharding.assign_units(units, world, 3_000_000)
        assert sorted(i for r in plan for i in r) == list(range(48))
        for mine in plan:
            assert len(mine) == 48 // world
            view_ids = [units[i].view for i in mine]
            settings = [dgr.GaussianRasterizationSettings(units[i].H, units[i].W, 0.45, 0.3, torch.zeros(3), 1.0, torch.eye(4),
                                                          torch.eye(4), 3, torch.zeros(3), False, False, True, True, 1.0)
                        for i in mine]
            groups = dgr._view_groups(settings, view_ids=view_ids)
            assert len(groups) == 8 // world and all(len(g) == 6 for g in groups)
            for g in groups:
                assert len({view_ids[k] for k in g}) == 1
                assert [units[mine[k]].scale for k in g] == [1, 2, 4, 8, 16, 32]


def test_gradient_sink_plumbing_cpu(monkeypatch):
    """`set_gradient_sink` host logic without a GPU (the kernels are replaced by stand-ins that honour `out=` the way
    `_C.rasterize_gaussians_backward` does): validation errors, the autograd Function hands the sink's views to the
    backward, autograd adopts them as .grad (no copy), `gradient_sink_installed` reports it, and None switches it off."""
    from harness import msgs_state
    dgr = msgs_state.import_product()
    P, M = 7, 16
    x = sh.GradientExchange(P, "cpu")
    with pytest.raises(Exception):
        dgr.set_gradient_sink({"mean3D": x.shaped()["mean3D"]})
    bad = dict(x.shaped())
    bad["sh"] = bad["sh"].double()
    with pytest.raises(Exception):
        dgr.set_gradient_sink(bad)

    def fake_fwd(*a):
        z = torch.zeros(0, dtype=torch.uint8)
        return (1, torch.zeros(3, 2, 2), torch.zeros(1, 2, 2), torch.zeros(1, 2, 2), torch.ones(P, dtype=torch.int32),
                torch.zeros(P), z, z, z)

    seen = {}

    def fake_bwd(*a, skip=(), out=None):
        seen["out"] = out
        shapes = {"mean3D": (P, 3), "sh": (P, M, 3), "opacity": (P, 1), "scale": (P, 3), "rot": (P, 4)}
        g = {k: (out[k].view(s) if out else torch.empty(s)) for k, s in shapes.items()}
        for i, k in enumerate(shapes):
            g[k].fill_(float(i + 1))
        return (torch.zeros(P, 3), None, g["opacity"], None, None, g["mean3D"], None, g["sh"], g["scale"], g["rot"])

    monkeypatch.setattr(dgr._C, "rasterize_gaussians", fake_fwd)
    monkeypatch.setattr(dgr._C, "rasterize_gaussians_backward", fake_bwd)
    eye = torch.eye(4)
    rs = dgr.GaussianRasterizationSettings(2, 2, 1.0, 1.0, torch.zeros(3), 1.0, eye, eye, 3, torch.zeros(3), False, False,
                                           False, False, 1.0)

    def step():
        p = dict(means3D=torch.zeros(P, 3), shs=torch.zeros(P, M, 3), opacities=torch.zeros(P, 1),
                 scales=torch.ones(P, 3), rotations=torch.ones(P, 4))
        for v in p.values():
            v.requires_grad_(True)
        m2 = torch.zeros(P, 3, requires_grad=True)
        color = dgr.GaussianRasterizer(rs)(
            means3D=p["means3D"], means2D=m2, opacities=p["opacities"], occ_multiplier=torch.ones(P, 4, 1),
            dc_delta=torch.zeros(P, 12, 1), shs=p["shs"], base_mask=torch.ones(P, dtype=torch.bool),
            scales=p["scales"], rotations=p["rotations"], max_pixel_sizes=torch.zeros(P), min_pixel_sizes=torch.zeros(P))[0]
        color.sum().backward()
        return [p[k] for k in ("means3D", "shs", "opacities", "scales", "rotations")]

    try:
        dgr.set_gradient_sink(x.shaped())
        ps = step()
        assert seen["out"] is not None and set(seen["out"]) == {"mean3D", "sh", "opacity", "scale", "rot"}
        assert dgr.gradient_sink_installed(*ps)
        assert ps[1].grad.shape == (P, M, 3) and ps[1].grad.data_ptr() == x.views["sh"].data_ptr()
        assert float(x.views["sh"][0]) == 2.0 and float(x.views["rot"][-1]) == 5.0
        assert float(x.flat.min()) >= 1.0                      # the five gradients tile the whole flat buffer
        dgr.set_gradient_sink(None)
        ps = step()
        assert seen["out"] is None and not dgr.gradient_sink_installed(*ps)
        # a sink of the wrong size fails loudly in the backward
        dgr.set_gradient_sink(sh.GradientExchange(P + 1, "cpu").shaped())
        with pytest.raises(Exception):
            step()
    finally:
        dgr.set_gradient_sink(None)


def _sink_worker(rank, world, port, P, M, out_dir):
    """One rank of a data-parallel step through the autograd API with the gradient sink: the (stand-in) backward writes this
    rank's gradients into the flat buffer, ONE all_reduce of the flat buffer follows, the parameters' .grad hold the sum."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from harness import msgs_state
        dgr = msgs_state.import_product()
        shapes = {"mean3D": (P, 3), "sh": (P, M, 3), "opacity": (P, 1), "scale": (P, 3), "rot": (P, 4)}

        def fake_fwd(*a):
            z = torch.zeros(0, dtype=torch.uint8)
            return (1, torch.zeros(3, 2, 2), torch.zeros(1, 2, 2), torch.zeros(1, 2, 2), torch.ones(P, dtype=torch.int32),
                    torch.zeros(P), z, z, z)

        def fake_bwd(*a, skip=(), out=None):
            gen = torch.Generator().manual_seed(77 + rank)
            g = {}
            for k, s in shapes.items():
                g[k] = out[k].view(s) if out else torch.empty(s)
                g[k].copy_(torch.randn(s, generator=gen))
            return (torch.zeros(P, 3), None, g["opacity"], None, None, g["mean3D"], None, g["sh"], g["scale"], g["rot"])

        dgr._C.rasterize_gaussians, dgr._C.rasterize_gaussians_backward = fake_fwd, fake_bwd
        eye = torch.eye(4)
        rs = dgr.GaussianRasterizationSettings(2, 2, 1.0, 1.0, torch.zeros(3), 1.0, eye, eye, 3, torch.zeros(3), False,
                                               False, False, False, 1.0)
        x = sh.GradientExchange(P, "cpu")
        dgr.set_gradient_sink(x.shaped())
        p = dict(means3D=torch.zeros(P, 3), shs=torch.zeros(P, M, 3), opacities=torch.zeros(P, 1),
                 scales=torch.ones(P, 3), rotations=torch.ones(P, 4))
        for v in p.values():
            v.requires_grad_(True)
        color = dgr.GaussianRasterizer(rs)(
            means3D=p["means3D"], means2D=torch.zeros(P, 3, requires_grad=True), opacities=p["opacities"],
            occ_multiplier=torch.ones(P, 4, 1), dc_delta=torch.zeros(P, 12, 1), shs=p["shs"],
            base_mask=torch.ones(P, dtype=torch.bool), scales=p["scales"], rotations=p["rotations"],
            max_pixel_sizes=torch.zeros(P), min_pixel_sizes=torch.zeros(P))[0]
        color.sum().backward()
        order = ("means3D", "shs", "opacities", "scales", "rotations")
        assert dgr.gradient_sink_installed(*(p[k] for k in order))
        dist.all_reduce(x.flat)                                  # the step's ONE collective
        if rank == 0:
            torch.save({k: p[k].grad.clone() for k in order}, os.path.join(out_dir, "sink_rank0.pt"))
    finally:
        dist.destroy_process_group()


def test_gradient_sink_one_collective_world_size_2(tmp_path):
    P, M, world = 33, 16, 2
    mp.spawn(_sink_worker, args=(world, _free_port(), P, M, str(tmp_path)), nprocs=world, join=True)
    got = torch.load(os.path.join(str(tmp_path), "sink_rank0.pt"))
    shapes = {"mean3D": (P, 3), "sh": (P, M, 3), "opacity": (P, 1), "scale": (P, 3), "rot": (P, 4)}
    want = {k: torch.zeros(s) for k, s in shapes.items()}
    for rank in range(world):
        gen = torch.Generator().manual_seed(77 + rank)
        for k, s in shapes.items():
            want[k] += torch.randn(s, generator=gen)
    for pk, sk in (("means3D", "mean3D"), ("shs", "sh"), ("opacities", "opacity"), ("scales", "scale"), ("rotations", "rot")):
        assert torch.allclose(got[pk], want[sk], rtol=1e-6, atol=1e-6), pk
