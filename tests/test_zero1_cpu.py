"""ZeRO-1 over a real DP group (gloo): parity with un-sharded AdamW + clipping, checkpoint conversion between DP sizes,
DCP save/load (reference ``test/integration/zero1/test_zero1.py``, ``convert_checkpoints/run_zero1_test.sh``,
``zero1_dcp/test_zero1_dcp.py``)."""
from __future__ import annotations

import os
import tempfile

import torch

from conftest import run_distributed


def _net(seed=0):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(24, 40), torch.nn.Tanh(), torch.nn.Linear(40, 8))


def _dp_worker(rank, world, tmp):
    import torch.distributed as dist

    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.optimizer import NeuronZero1Optimizer, convert_zero_checkpoints as czc
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    cfg = nxd.neuronx_distributed_config(optimizer_config={"zero_one_enabled": True, "grad_clipping": True,
                                                           "max_grad_norm": 0.5})
    assert ps.get_data_parallel_size() == world
    model = nxd.initialize_parallel_model(cfg, _net, device=torch.device("cpu"))
    opt = nxd.initialize_parallel_optimizer(cfg, torch.optim.AdamW, model.parameters(), lr=1e-2, weight_decay=0.05)
    assert isinstance(opt.optimizer, NeuronZero1Optimizer) and opt.optimizer.buckets[0].dp == world
    ref = _net()
    ref_opt = torch.optim.AdamW(ref.parameters(), lr=1e-2, weight_decay=0.05)
    g = torch.Generator().manual_seed(5)
    for step in range(4):
        xs = torch.randn(world, 6, 24, generator=g)
        ys = torch.randn(world, 6, 8, generator=g)
        opt.zero_grad()
        ((model(xs[rank]) - ys[rank]) ** 2).mean().backward()
        opt.step()
        ref_opt.zero_grad()
        sum(((ref(xs[r]) - ys[r]) ** 2).mean() for r in range(world)).div(world).backward()
        norm = torch.nn.utils.clip_grad_norm_(ref.parameters(), 0.5)
        ref_opt.step()
        assert abs(float(opt.grad_norm) - float(norm)) < 1e-4 * max(1.0, float(norm))
        for p, q in zip(model.parameters(), ref.parameters()):
            torch.testing.assert_close(p, q, atol=1e-5, rtol=1e-4)

    # sharded save -> convert to full -> re-shard to another DP size -> merge again: identical tensors
    d = os.path.join(tmp, "ck")
    nxd.save_checkpoint(d, "t", model=model, optimizer=opt)
    optim_dir = os.path.join(d, "t", "optim")
    assert os.path.exists(os.path.join(optim_dir, f"dp_rank_{rank:02d}_tp_rank_00_pp_rank_00.pt"))
    if rank == 0:
        czc.convert(optim_dir, os.path.join(tmp, "full"), convert_to_full=True, num_workers=1)
        full = torch.load(os.path.join(tmp, "full", "dp_rank_00_tp_rank_00_pp_rank_00.pt"), weights_only=False)
        w0 = dict(ref.named_parameters())["0.weight"]
        torch.testing.assert_close(full["base_state"][0], w0.detach(), atol=1e-5, rtol=1e-4)
        assert full["state"][0]["exp_avg"].shape == w0.shape
        czc.convert(os.path.join(tmp, "full"), os.path.join(tmp, "dp3"), convert_to_sharded=True, dp_size=3, num_workers=1)
        shards = [torch.load(os.path.join(tmp, "dp3", f"dp_rank_{r:02d}_tp_rank_00_pp_rank_00.pt"), weights_only=False)
                  for r in range(3)]
        again = czc.merge_shards(shards)
        for i in full["base_state"]:
            assert torch.equal(again["base_state"][i], full["base_state"][i])
            assert torch.equal(again["state"][i]["exp_avg_sq"], full["state"][i]["exp_avg_sq"])
        # the same round trip through the reference's helper names (args namespace, <ckpt>/<tag> as input_dir)
        from types import SimpleNamespace
        a = SimpleNamespace(input_dir=os.path.join(d, "t"), output_dir=os.path.join(tmp, "dp4"), dp_size=4)
        assert not czc.is_full(a) and not czc.is_xser(a) and czc.get_parallel_info(a) == (world, 1, 1)
        assert czc.is_full(SimpleNamespace(input_dir=os.path.join(tmp, "full")))
        merged = czc.merge_optim_dp_checkpoints(a, 0, 0)
        assert len(czc.split_and_save_ckpts(a, merged, 0, 0)) == 4
        again4 = czc.merge_shards([torch.load(os.path.join(tmp, "dp4", f"dp_rank_{r:02d}_tp_rank_00_pp_rank_00.pt"),
                                              weights_only=False) for r in range(4)])
        for i in full["base_state"]:
            assert torch.equal(again4["base_state"][i], full["base_state"][i])
    dist.barrier()

    # DCP variant: one optim/ directory with .metadata, every rank's tensors under rank-suffixed names
    d2 = os.path.join(tmp, "ck_dcp")
    before = opt.state_dict()
    nxd.save_checkpoint(d2, "t", model=model, optimizer=opt, use_zero1_dcp=True)
    assert os.path.exists(os.path.join(d2, "t", "optim", ".metadata"))
    for b in opt.optimizer.buckets:
        b.master.zero_()
        b.exp_avg.zero_()
    nxd.load_checkpoint(d2, "t", model=model, optimizer=opt, use_zero1_dcp=True)
    after = opt.state_dict()
    for i in before["base_state"]:
        assert torch.equal(before["base_state"][i], after["base_state"][i])
        if i in before["state"]:
            assert torch.equal(before["state"][i]["exp_avg"], after["state"][i]["exp_avg"])


def _dcp_async_worker(rank, world, tmp):
    """Asynchronous save with the DCP optimizer path (reference unit test ``test_checkpoint.py:197-332``): the writer
    thread must produce ``optim/.metadata`` + one ``__<rank>_0.distcp`` per rank without issuing a collective, and the
    state must round-trip."""
    import torch.distributed as dist

    from neuronx_distributed_llama3_2_b200 import trainer as nxd

    cfg = nxd.neuronx_distributed_config(optimizer_config={"zero_one_enabled": True, "grad_clipping": True,
                                                           "max_grad_norm": 1.0})
    model = nxd.initialize_parallel_model(cfg, _net, device=torch.device("cpu"))
    opt = nxd.initialize_parallel_optimizer(cfg, torch.optim.AdamW, model.parameters(), lr=1e-2)
    g = torch.Generator().manual_seed(11 + rank)
    for _ in range(2):
        opt.zero_grad()
        (model(torch.randn(6, 24, generator=g)) ** 2).mean().backward()
        opt.step()
    before = opt.state_dict()
    calls = []
    real = {n: getattr(dist, n) for n in ("all_reduce", "all_gather", "all_gather_object", "gather_object", "broadcast",
                                          "broadcast_object_list", "scatter_object_list", "reduce_scatter_tensor")}

    def spy(name):
        def f(*a, **k):
            import threading

            if threading.current_thread() is not threading.main_thread():
                calls.append(name)
            return real[name](*a, **k)
        return f

    for n in real:
        setattr(dist, n, spy(n))
    try:
        d = os.path.join(tmp, "ck_async")
        nxd.save_checkpoint(d, "t", model=model, optimizer=opt, use_zero1_dcp=True, async_save=True)
        nxd.finalize_checkpoint()
    finally:
        for n, f in real.items():
            setattr(dist, n, f)
    assert not calls, f"collectives issued from the checkpoint writer thread: {calls}"
    names = set(os.listdir(os.path.join(d, "t", "optim")))
    assert ".metadata" in names and f"__{rank}_0.distcp" in names, names
    assert not [n for n in names if n.startswith("write_results")], names
    assert os.path.exists(os.path.join(d, "t", "done"))
    for b in opt.optimizer.buckets:
        b.master.zero_()
        b.exp_avg.zero_()
        b.exp_avg_sq.zero_()
    nxd.load_checkpoint(d, "t", model=model, optimizer=opt, use_zero1_dcp=True)
    after = opt.state_dict()
    for i in before["base_state"]:
        assert torch.equal(before["base_state"][i], after["base_state"][i])
        if i in before["state"]:
            assert torch.equal(before["state"][i]["exp_avg_sq"], after["state"][i]["exp_avg_sq"])


def test_zero1_dcp_async_save_is_collective_free():
    with tempfile.TemporaryDirectory() as tmp:
        run_distributed(_dcp_async_worker, 2, tmp)


def test_zero1_dp2_parity_convert_and_dcp():
    with tempfile.TemporaryDirectory() as tmp:
        run_distributed(_dp_worker, 2, tmp)
