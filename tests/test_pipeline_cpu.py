"""NxDPPModel on CPU/gloo: PP=2 (1F1B) and PP=2 x virtual 2 (interleaved) against the un-partitioned model --
loss, grads (incl. the tied embedding living on both stages), PP-agnostic state dict names, trainer + ZeRO-1 + checkpoint."""
from __future__ import annotations

import os
import tempfile

import pytest
import torch

from conftest import run_distributed


def _build(num_layers=4):
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

    torch.manual_seed(0)
    cfg = tiny_llama(num_hidden_layers=num_layers, sequence_parallel_enabled=False)
    return cfg, LlamaForCausalLM(cfg)


def _pp_worker(rank, world, vp, tmp):
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models.llama import LlamaDecoderLayer
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    nxd_cfg = nxd.neuronx_distributed_config(
        tensor_parallel_size=1, pipeline_parallel_size=2,
        pipeline_config={"transformer_layer_cls": LlamaDecoderLayer, "num_microbatches": 4, "auto_partition": True,
                         "input_names": ["input_ids", "labels"], "virtual_pipeline_size": vp},
        optimizer_config={"zero_one_enabled": True, "grad_clipping": True, "max_grad_norm": 1.0})
    assert ps.get_pipeline_model_parallel_size() == 2
    cfg, ref = _build()
    ids = torch.randint(0, cfg.vocab_size, (8, 32), generator=torch.Generator().manual_seed(3))
    # oracle: mean over 4 micro-batches of 2 sequences each == what the pipeline computes
    ref_losses = [ref(c, labels=c) for c in torch.chunk(ids, 4, dim=0)]
    ref_loss = torch.stack(ref_losses).mean()
    ref_loss.backward()
    ref_grads = {n: p.grad.clone() for n, p in ref.named_parameters()}

    model = nxd.initialize_parallel_model(nxd_cfg, lambda: _build()[1], device=torch.device("cpu"))
    pp = model.module
    assert pp.num_stages == 2 * vp and len(pp.local_stage_ids) == vp
    loss = model.run_train(input_ids=ids, labels=ids)
    if rank == world - 1:
        assert abs(loss.item() - ref_loss.item()) < 1e-5, (loss.item(), ref_loss.item())
    else:
        assert loss is None
    names = []
    for n, p in model.named_parameters():
        names.append(n)
        rn = "model.embed_tokens.weight" if n == "lm_head.weight" else n   # tied: one tensor, two original names
        assert rn in ref_grads, f"local parameter name {n} is not an original name"
        assert p.grad is not None, n
        torch.testing.assert_close(p.grad, ref_grads[rn], atol=2e-5, rtol=1e-3, msg=lambda m: f"{n}: {m}")
    # the tied embedding is owned by the first stage (embed) and the last stage (lm_head): both see the SUM
    assert any(n.endswith("embed_tokens.weight") or n.endswith("lm_head.weight") for n in names)
    all_names = [None] * world
    torch.distributed.all_gather_object(all_names, names)
    covered = set(all_names[0]) | set(all_names[1])
    assert covered >= {n for n in ref_grads if not n.startswith("lm_head")}, set(ref_grads) - covered

    # optimizer + checkpoint through the trainer API
    opt = nxd.initialize_parallel_optimizer(nxd_cfg, torch.optim.AdamW, model.parameters(), lr=1e-2)
    for p in model.parameters():
        p.grad = None
    opt.zero_grad()
    l0 = model.run_train(input_ids=ids, labels=ids)
    opt.step()
    opt.zero_grad()
    l1 = model.run_train(input_ids=ids, labels=ids)
    if rank == world - 1:
        assert l1.item() < l0.item()
    d = os.path.join(tmp, f"pp_ckpt_vp{vp}")
    nxd.save_checkpoint(d, "s1", model=model, optimizer=opt)
    assert os.path.exists(os.path.join(d, "s1", "model", f"dp_rank_00_tp_rank_00_pp_rank_{rank:02d}.pt"))
    before = {k: v.clone() for k, v in model.state_dict().items()}
    with torch.no_grad():
        for p in model.parameters():
            p.mul_(0.5)
    nxd.load_checkpoint(d, "s1", model=model, optimizer=opt)
    for k, v in model.state_dict().items():
        assert torch.equal(v, before[k]), k
    # eval path
    outs = model.run_eval(input_ids=ids, labels=ids)
    if rank == world - 1:
        assert len(outs) == 4


@pytest.mark.parametrize("vp", [1, 2])
def test_pp2_matches_single_process(vp):
    with tempfile.TemporaryDirectory() as tmp:
        run_distributed(_pp_worker, 2, vp, tmp)


def _tp_pp_worker(rank, world):
    """BASELINE.json config #3 in miniature: TP=2 x PP=2 with sequence parallelism + activation checkpointing."""
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama
    from neuronx_distributed_llama3_2_b200.models.llama import LlamaDecoderLayer
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    nxd_cfg = nxd.neuronx_distributed_config(
        tensor_parallel_size=2, pipeline_parallel_size=2, sequence_parallel=True,
        pipeline_config={"transformer_layer_cls": LlamaDecoderLayer, "num_microbatches": 2, "auto_partition": True,
                         "input_names": ["input_ids", "labels"], "broadcast_and_average_loss": True},
        optimizer_config={"zero_one_enabled": True, "grad_clipping": True, "max_grad_norm": 1.0},
        activation_checkpoint_config="full")
    assert (ps.get_tensor_model_parallel_size(), ps.get_pipeline_model_parallel_size()) == (2, 2)
    torch.manual_seed(0)
    model = nxd.initialize_parallel_model(
        nxd_cfg, lambda: LlamaForCausalLM(tiny_llama(num_hidden_layers=4)), device=torch.device("cpu"))
    opt = nxd.initialize_parallel_optimizer(nxd_cfg, torch.optim.AdamW, model.parameters(), lr=1e-2)
    ids = torch.randint(0, 512, (4, 64), generator=torch.Generator().manual_seed(5))
    losses = []
    for _ in range(6):
        opt.zero_grad()
        losses.append(float(model.run_train(input_ids=ids, labels=ids)))
        opt.step()
    assert losses[-1] < losses[0] - 0.3, losses
    gathered = [None] * world
    torch.distributed.all_gather_object(gathered, losses)
    assert all(abs(g[-1] - losses[-1]) < 1e-5 for g in gathered), gathered  # broadcast_and_average_loss


def test_tp2_pp2_sp_activation_checkpointing_trains():
    run_distributed(_tp_pp_worker, 4)


def test_debug_mode_partition_and_names():
    from neuronx_distributed_llama3_2_b200.models.llama import LlamaDecoderLayer
    from neuronx_distributed_llama3_2_b200.pipeline import NxDPPModel
    from neuronx_distributed_llama3_2_b200.pipeline.partition import create_partitions

    assert create_partitions(4, 8) == [1, 3, 5]
    assert create_partitions(3, 8) == [1, 4]      # remainder goes to the later stages: sizes 2, 3, 3
    from neuronx_distributed_llama3_2_b200.pipeline.partition import (PipelineIO, adding_live_obj_for_previous_stages,
                                                                      iterate_graph_model_outputs)
    assert list(iterate_graph_model_outputs(("a",))) == ["a"] and list(iterate_graph_model_outputs((("a", "b"),))) == ["a", "b"]
    ins, outs = [{"x": PipelineIO("x")}, {"h0": PipelineIO("h0")}, {"h1": PipelineIO("h1")}], [{"h0": PipelineIO("h0")},
                                                                                                {"h1": PipelineIO("h1")}, {}]
    adding_live_obj_for_previous_stages(ins, outs, "x", 2)      # stage 3 wants the model input-derived value "x"
    assert "x" in outs[0] and "x" in ins[1] and "x" in outs[1] and "x" in ins[2] and "x" in outs[2]
    with pytest.raises(RuntimeError):
        adding_live_obj_for_previous_stages(ins, outs, "nope", 2)
    cfg, m = _build(num_layers=8)
    pp = NxDPPModel(m, transformer_layer_cls=LlamaDecoderLayer, num_microbatches=2, input_names=["input_ids", "labels"],
                    auto_partition=True, _debug_mode=True, _debug_pp_size=4, _debug_pp_rank=1)
    assert pp.pipeline_cuts == ["model.layers.1", "model.layers.3", "model.layers.5"]
    local = [n for n, _ in pp.local_named_parameters()]
    assert all(n.startswith("model.layers.2.") or n.startswith("model.layers.3.") for n in local), local
    assert set(pp.local_state_dict().keys()) == set(local)
    # parameters of other stages were released
    assert m.model.layers[0].mlp.down_proj.weight.numel() == 0
    with pytest.raises(RuntimeError):
        pp(torch.zeros(1))
    # reference-named helpers
    layers = pp.get_model_layers()
    assert len(layers) == 2 and all(isinstance(l, LlamaDecoderLayer) for l in layers)
    assert pp.get_current_stage(0) == 1 and not pp.is_last_pp_rank_last_model_chunk(0)
    sd = pp.local_state_dict()
    loc = pp.translate_origin_state_dict_to_local_state_dict(dict(sd, **{"model.layers.7.mlp.down_proj.weight": torch.zeros(1)}))
    assert loc and all(k.startswith("submod_1.") for k in loc) and len(loc) == len(sd)
    back = pp.translate_local_state_dict_to_origin_state_dict(loc)
    assert set(back) == set(sd)
    chunks = pp.construct_state_dict_per_model_chunk(sd)
    assert len(chunks) == 1 and len(chunks[0]) == len(sd)
    mbs = list(pp.get_batch_iterator({"input_ids": torch.zeros(4, 8, dtype=torch.long), "flag": True}))
    assert len(mbs) == 2 and mbs[0]["input_ids"].shape == (2, 8) and mbs[1]["flag"] is True
    pp.clear_minibatch_state()
    assert pp.stage_outputs == {} and pp.losses == []
    a = torch.randn(3, requires_grad=True)
    out = a * 2
    NxDPPModel.custom_backward([out], [torch.ones(3)])
    assert torch.equal(a.grad, torch.full((3,), 2.0))
    pp.move_model_to_device(torch.device("cpu"))


class _Inner(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.linear1, self.linear2, self.linear3 = (torch.nn.Linear(2, 2) for _ in range(3))

    def forward(self, x):
        return (self.linear3(self.linear2(self.linear1(x))) + x).transpose(0, 1)


class _Nested(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.my_mod = _Inner()
        self.linear4 = torch.nn.Linear(2, 2)

    def forward(self, x):
        return self.linear4(self.my_mod(x))


from neuronx_distributed_llama3_2_b200.parallel.loss_functions import parallel_cross_entropy  # noqa: E402  (module level:
# torch.fx auto-wraps a function by patching it in the globals of the module that calls it)


class _WithParallel(torch.nn.Module):
    def __init__(self):
        super().__init__()
        from neuronx_distributed_llama3_2_b200.parallel.layers import ColumnParallelLinear

        self.my_mod = _Inner()
        self.cpl = ColumnParallelLinear(2, 6, gather_output=False)

    def forward(self, x, labels):
        return parallel_cross_entropy(self.cpl(self.my_mod(x)), labels)


def test_trace_helpers_leaf_names_and_tracer_selection():
    """The reference's ``test/unit_test/pipeline/test_trace.py`` cases: concrete args, tracer selection by keyword,
    node sequence of a traced module, leaves given by class NAME, auto-wrapped parallel functions."""
    from neuronx_distributed_llama3_2_b200.pipeline import trace as T

    class Optional_(torch.nn.Module):
        def forward(self, x, unused=None, flag=True):
            return x

    assert T.get_concrete_args(Optional_(), ["x"]) == {"unused": None, "flag": True}
    assert T.get_concrete_args(_Nested(), ["x"]) == {}
    assert T.get_tracer_class(_Inner()) is T.TorchTracerWrapper
    assert T.get_tracer_class(_Inner(), "hf") is T.HFTracerWrapper
    assert T.get_tracer_class(_Inner(), "torch") is T.TorchTracerWrapper
    with pytest.raises(ValueError):
        T.get_tracer_class(_Inner(), "test")
    ops = [(n.op, n.name) for n in T.trace_model(_Inner(), input_names=["x"]).graph.nodes]
    assert ops == [("placeholder", "x"), ("call_module", "linear1"), ("call_module", "linear2"), ("call_module", "linear3"),
                   ("call_function", "add"), ("call_method", "transpose"), ("output", "output")]
    ops = [(n.op, n.name) for n in T.trace_model(_Nested(), input_names=["x"], leaf_modules=["_Inner"]).graph.nodes]
    assert ops == [("placeholder", "x"), ("call_module", "my_mod"), ("call_module", "linear4"), ("output", "output")]

    gm = T.trace_model(_WithParallel(), input_names=["x", "labels"], leaf_modules=[_Inner])
    ops = [(n.op, n.name) for n in gm.graph.nodes]
    assert ops == [("placeholder", "x"), ("placeholder", "labels"), ("call_module", "my_mod"), ("call_module", "cpl"),
                   ("call_function", "parallel_cross_entropy"), ("output", "output")]
    x, y = torch.randn(2, 2), torch.randint(0, 6, (2,))
    torch.testing.assert_close(gm(x, y), parallel_cross_entropy(gm.cpl(gm.my_mod(x)), y))


def test_partition_io_maps_and_shared_weights_of_eight_stage_model():
    """The reference's fixture (``test/unit_test/pipeline/test_base.py``): an 8-block decoder with tied embeddings cut after
    every block but the last -> 8 stages; hidden states hop stage to stage, ``labels`` is consumed where the loss lives,
    the tied weight is reported as shared between stage 0 and stage 7."""
    from neuronx_distributed_llama3_2_b200.pipeline import partition as P
    from neuronx_distributed_llama3_2_b200.pipeline.trace import trace_model

    class Block(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.ln, self.fc = torch.nn.LayerNorm(16), torch.nn.Linear(16, 16)

        def forward(self, h, mask):
            return h + self.fc(self.ln(h)) * mask

    class GPT(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.wte = torch.nn.Embedding(50, 16)
            self.h = torch.nn.ModuleList(Block() for _ in range(8))
            self.ln_f = torch.nn.LayerNorm(16)
            self.lm_head = torch.nn.Linear(16, 50, bias=False)
            self.lm_head.weight = self.wte.weight

        def forward(self, input_ids, attention_mask, labels):
            h = self.wte(input_ids)
            m = attention_mask.unsqueeze(-1).to(h.dtype)
            for blk in self.h:
                h = blk(h, m)
            logits = self.lm_head(self.ln_f(h))
            return torch.nn.functional.cross_entropy(logits.flatten(0, 1), labels.flatten())

    torch.manual_seed(0)
    model = GPT()
    traced = trace_model(model, input_names=["input_ids", "attention_mask", "labels"], leaf_modules=["Block"])
    cuts = [f"h.{i}" for i in range(7)]
    split = P.partition_traced_model(traced, cuts)
    stages = [n for n, _ in split.named_children()]
    assert stages == [f"submod_{i}" for i in range(8)]
    inputs, ios = P.analyze_pipeline_module(split)
    assert inputs == ["input_ids", "attention_mask", "labels"]
    assert [n for n, _ in ios[0].model_inputs] == ["input_ids", "attention_mask"] and ios[0].recv == []
    assert [n for n, _ in ios[7].model_inputs] == ["labels"] and ios[7].send == [] and ios[7].returns_model_output
    for k in range(1, 8):
        # two values cross every boundary: the hidden state and the broadcast mask computed on stage 0
        assert len(ios[k].recv) == 2 and {io.name for io in ios[k].recv} == {io.name for io in ios[k - 1].send}
        assert all(io.input_idx is not None for io in ios[k].recv)
    shared = P.analyze_shared_weights_across_stages(split)
    assert len(shared) == 1 and sorted(s for _, s in shared[0]) == [0, 7]
    assert {n for n, _ in shared[0]} == {"wte.weight", "lm_head.weight"}
    ids = torch.randint(0, 50, (2, 6))
    mask, labels = torch.ones(2, 6), torch.randint(0, 50, (2, 6))
    torch.testing.assert_close(split(ids, mask, labels), model(ids, mask, labels))
    with pytest.raises(ValueError):
        P.partition_traced_model(traced, ["h.42"])


def _pp_mixtral_worker(rank, world):
    """Mixtral under pipeline parallelism: the decoder layer returns (hidden, residual, router logits); the router logits
    of the first stage travel to the last one, where the load-balancing loss is one opaque node of the traced graph."""
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models import MixtralForCausalLM, tiny_mixtral
    from neuronx_distributed_llama3_2_b200.models.mixtral import MixtralDecoderLayer

    nxd_cfg = nxd.neuronx_distributed_config(
        tensor_parallel_size=1, pipeline_parallel_size=2,
        pipeline_config={"transformer_layer_cls": MixtralDecoderLayer, "num_microbatches": 2, "auto_partition": True,
                         "input_names": ["input_ids", "labels"]},
        optimizer_config={"zero_one_enabled": True, "grad_clipping": True, "max_grad_norm": 1.0})

    def build():
        torch.manual_seed(0)
        return MixtralForCausalLM(tiny_mixtral(sequence_parallel_enabled=False, num_hidden_layers=4, capacity_factor=1.0))

    ref = build()
    ids = torch.randint(0, 512, (4, 32), generator=torch.Generator().manual_seed(1))
    ref_loss = torch.stack([ref(c, labels=c) for c in torch.chunk(ids, 2, dim=0)]).mean()
    ref_loss.backward()
    ref_grads = {n: p.grad.clone() for n, p in ref.named_parameters() if p.grad is not None}
    model = nxd.initialize_parallel_model(nxd_cfg, build, device=torch.device("cpu"))
    loss = model.run_train(input_ids=ids, labels=ids)
    if rank == world - 1:
        assert abs(float(loss) - float(ref_loss)) < 1e-5, (float(loss), float(ref_loss))
    checked = 0
    for n, p in model.named_parameters():
        if p.grad is not None and n in ref_grads:
            torch.testing.assert_close(p.grad, ref_grads[n], atol=2e-5, rtol=1e-3, msg=lambda m: f"{n}: {m}")
            checked += 1
    assert checked > 5
    # the aux loss reaches the routers of the FIRST stage too (their logits are inputs of the loss on the last stage)
    if rank == 0:
        g = dict(model.named_parameters())["layers.0.mlp.router.linear_router.weight"].grad
        assert g is not None and float(g.abs().sum()) > 0
    opt = nxd.initialize_parallel_optimizer(nxd_cfg, torch.optim.AdamW, model.parameters(), lr=1e-2)
    losses = []
    for _ in range(3):
        opt.zero_grad()
        out = model.run_train(input_ids=ids, labels=ids)
        opt.step()
        if out is not None:
            losses.append(float(out))
    if rank == world - 1:
        assert losses[-1] < losses[0]


def test_pp2_mixtral_matches_single_process():
    run_distributed(_pp_mixtral_worker, 2)


def _pp_family_worker(rank, world, fam):
    """GPT-NeoX, DBRX and BERT under PP = 2.  DBRX's final norm is traced through and BERT adds ``self.decoder_bias`` in
    its own forward: such root-level parameters are handed to the stage that reads them and keep their original names."""
    from neuronx_distributed_llama3_2_b200 import models as M
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models.bert import BertLayer
    from neuronx_distributed_llama3_2_b200.models.dbrx import DbrxDecoderLayer
    from neuronx_distributed_llama3_2_b200.models.gpt_neox import GPTNeoXLayer

    layer_cls = {"neox": GPTNeoXLayer, "dbrx": DbrxDecoderLayer, "bert": BertLayer}[fam]
    names = ["input_ids", "labels"] + (["next_sentence_label"] if fam == "bert" else [])
    nxd_cfg = nxd.neuronx_distributed_config(
        tensor_parallel_size=1, pipeline_parallel_size=2,
        pipeline_config={"transformer_layer_cls": layer_cls, "num_microbatches": 2, "auto_partition": True,
                         "input_names": names},
        optimizer_config={"zero_one_enabled": True, "grad_clipping": True, "max_grad_norm": 1.0})

    def build():
        torch.manual_seed(0)
        if fam == "neox":
            return M.GPTNeoXForCausalLM(M.tiny_gpt_neox(num_hidden_layers=4))
        if fam == "dbrx":
            return M.DbrxForCausalLM(M.tiny_dbrx(sequence_parallel_enabled=False, num_hidden_layers=4, capacity_factor=1.0))
        return M.BertForPreTraining(M.tiny_bert(num_hidden_layers=4))

    ref = build()
    ids = torch.randint(0, 512, (4, 32), generator=torch.Generator().manual_seed(1))
    kw = dict(input_ids=ids, labels=ids)
    if fam == "bert":
        kw["next_sentence_label"] = torch.tensor([0, 1, 1, 0])
    ref_loss = torch.stack([ref(**{k: torch.chunk(v, 2, 0)[i] for k, v in kw.items()}) for i in range(2)]).mean()
    ref_loss.backward()
    # tied weights (BERT's MLM decoder = word embedding) have one tensor and two original names
    ref_grads = {n: p.grad for n, p in ref.named_parameters(remove_duplicate=False) if p.grad is not None}
    model = nxd.initialize_parallel_model(nxd_cfg, build, device=torch.device("cpu"))
    loss = model.run_train(**kw)
    if rank == world - 1:
        assert abs(float(loss.detach()) - float(ref_loss.detach())) < 1e-5, (fam, float(loss), float(ref_loss))
    local = dict(model.named_parameters())
    assert set(local) <= set(ref_grads), set(local) - set(ref_grads)
    for n, p in local.items():
        if n in ref_grads:
            assert p.grad is not None, n
            torch.testing.assert_close(p.grad, ref_grads[n], atol=2e-5, rtol=1e-3, msg=lambda m: f"{fam} {n}: {m}")
    root_level = {"dbrx": "norm.weight", "bert": "decoder_bias"}.get(fam)
    if root_level and rank == world - 1:
        assert root_level in local and local[root_level].grad is not None
    assert root_level is None or root_level in model.state_dict() or rank != world - 1


@pytest.mark.parametrize("fam", ["neox", "dbrx", "bert"])
def test_pp2_other_families_match_single_process(fam):
    run_distributed(_pp_family_worker, 2, fam)


class _Payload:
    def __init__(self, x):
        self.x = x

    def __eq__(self, other):
        return isinstance(other, _Payload) and self.x == other.x


def _comm_relay_worker(rank, world):
    """The reference's ``test/integration/parallel_layers/pipeline/test_comm.py``: a tensor and a python object are relayed
    stage by stage down the pipeline, then a gradient-channel tensor travels back up while activations still flow."""
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps
    from neuronx_distributed_llama3_2_b200.pipeline import comm
    from neuronx_distributed_llama3_2_b200.utils.serialization import TensorMeta

    ps.initialize_model_parallel(tensor_model_parallel_size=1, pipeline_model_parallel_size=world)
    comm.init_p2p_groups()
    pp, last = ps.get_pipeline_model_parallel_rank(), world - 1
    meta = TensorMeta(tensor_index=-1, dtype=torch.float32, shape=torch.Size([2, 3]), requires_grad=False, device=None)
    want = torch.arange(6, dtype=torch.float32).reshape(2, 3) * 1.5
    if pp == 0:
        comm.send(want.clone())
        comm.send_python_object({"step": 3, "obj": _Payload([1, 2])})
    else:
        got = comm.recv_from(meta)
        obj = comm.recv_python_object()
        assert torch.equal(got, want) and obj == {"step": 3, "obj": _Payload([1, 2])}
        if pp < last:
            comm.send(got)
            comm.send_python_object(obj)
    # both directions at once between neighbours: activation down (tag / communicator A), gradient up (B)
    if pp < last:
        comm.send(want + pp)
    if pp > 0:
        comm.send(want - pp, send_next=False)
    if pp > 0:
        assert torch.equal(comm.recv_from(meta), want + (pp - 1))
    if pp < last:
        assert torch.equal(comm.recv_from(meta, recv_prev=False), want - (pp + 1))
    comm.flush_sends()
    comm.reset_p2p_groups()


def test_pipeline_comm_relay_and_bidirectional_channels():
    run_distributed(_comm_relay_worker, 4)
