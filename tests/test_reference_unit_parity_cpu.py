"""The behaviours the reference's small unit tests pin down (``test/unit_test/{utils,wrapper,pipeline}``), restated against
this package: casting helpers, logger levels and rank gating, tensor serialisation for the pipeline engine, ``cumsum``,
NxD config defaults, post-partition hooks."""
from __future__ import annotations

import logging
from unittest.mock import MagicMock, patch

import pytest
import torch

import neuronx_distributed_llama3_2_b200 as nxd


# -- test_casting_utils.py -------------------------------------------------------------------------------------------------
def test_cast_if_autocast_enabled_and_verify():
    from neuronx_distributed_llama3_2_b200.parallel.utils import cast_if_autocast_enabled, verify_casted_dtype

    inputs = (torch.tensor([0.0]), torch.tensor([0.0]), 12)
    with torch.autocast(enabled=True, dtype=torch.bfloat16, device_type="cpu"):
        casted = cast_if_autocast_enabled(*inputs)
        assert all(x.dtype == torch.bfloat16 for x in casted if isinstance(x, torch.Tensor)) and casted[2] == 12
        verify_casted_dtype(casted)
        with pytest.raises(AssertionError, match="Datatype of tensor is expected to be torch.bfloat16, got torch.float32"):
            verify_casted_dtype(inputs)
        verify_casted_dtype({"k": [casted[0], (casted[1],)], "n": 3})
    with torch.autocast(enabled=False, device_type="cpu"):
        assert all(x.dtype == torch.float32 for x in cast_if_autocast_enabled(*inputs) if isinstance(x, torch.Tensor))
    verify_casted_dtype(inputs)                       # no autocast: anything goes


# -- test_logger.py ----------------------------------------------------------------------------------------------------------
def test_log_levels_and_rank_gating():
    from neuronx_distributed_llama3_2_b200.utils import logger as L

    levels = {"trace": logging.DEBUG, "debug": logging.DEBUG, "info": logging.INFO, "warning": logging.WARNING,
              "error": logging.ERROR, "fatal": logging.FATAL, "off": logging.FATAL + 1}
    try:
        for value, level in levels.items():
            with patch.dict("os.environ", {"NXD_LOG_LEVEL": value}):
                L.get_log_level.cache_clear()
                assert L.get_log_level() == level
        with patch.dict("os.environ", {"NXD_LOG_LEVEL": "unsupported"}):
            L.get_log_level.cache_clear()
            with pytest.raises(ValueError):
                L.get_log_level()
        with patch.dict("os.environ", {"NXD_LOG_LEVEL": "off"}):
            L.get_log_level.cache_clear()
            assert L.get_logger("nxd_test_off").disabled
    finally:
        L.get_log_level.cache_clear()
    fn = MagicMock()
    with patch("torch.distributed") as d:
        d.is_available.return_value = True
        d.is_initialized.return_value = True
        d.get_rank.return_value = 0
        assert L._rank0_only(fn)() == fn.return_value
        fn.assert_called_once_with()
        d.get_rank.return_value = 3
        other = MagicMock()
        assert L._rank0_only(other, default="skipped")() == "skipped" and not other.called
    with patch.dict("os.environ", {"RANK": "1"}):                 # before the process group exists: the launcher's RANK
        late = MagicMock()
        assert L._rank0_only(late)() is None and not late.called
    quiet = L.get_logger("nxd_test_gate", rank0_only=True)
    loud = L.get_logger("nxd_test_gate", rank0_only=False)
    with patch.dict("os.environ", {"RANK": "1"}), patch.object(quiet._inner, "info") as info:
        quiet.info("hidden on rank 1")
        assert not info.called
        loud.info("every rank")
        assert info.called


# -- test_serialization.py ----------------------------------------------------------------------------------------------------
def test_serialization_manager_round_trips():
    from neuronx_distributed_llama3_2_b200.utils.serialization import SerializationManager, TensorMeta

    class A:
        def __init__(self, x):
            self.x = x

    s = SerializationManager()
    a = A(3.2)
    ser, tx, meta = s.serialize(a)
    assert s.deserialize(ser, tx) == a and len(tx) == 0 and len(meta) == 0
    t = torch.ones(2, 3)
    ser, tx, meta = s.serialize(t)
    assert meta == [TensorMeta(tensor_index=0, dtype=torch.float32, shape=torch.Size([2, 3]), requires_grad=False,
                               device=torch.device("cpu"))]
    assert torch.equal(s.deserialize(ser, tx), t) and torch.equal(tx[0], t)
    data = {"a": 1, "b": [torch.ones(2, 4), torch.ones(2, 4), (1, 2)], "c": (1, 2, torch.tensor(1.0)),
            "d": torch.zeros(2, 4), "f": a}
    ser, tx, meta = s.serialize(data)
    assert [m.tensor_index for m in meta] == [0, 1, 2, 3] and [tuple(m.shape) for m in meta] == [(2, 4), (2, 4), (), (2, 4)]
    back = s.deserialize(ser, tx)
    assert back["a"] == 1 and back["b"][2] == (1, 2) and back["f"] == a and isinstance(back["c"], tuple)
    assert torch.equal(back["b"][1], data["b"][1]) and torch.equal(back["c"][2], data["c"][2])
    assert torch.equal(back["d"], data["d"])
    # the receiving stage rebuilds from ITS tensors (same structure, other values)
    swapped = s.deserialize(ser, [x + 1 for x in tx])
    assert torch.equal(swapped["d"], torch.ones(2, 4))


# -- test_tensor_utils.py ------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape", [(3, 2), (2, 3), (10, 10), (100, 2048), (2048, 100)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_cumsum_matches_torch(shape, dtype):
    from neuronx_distributed_llama3_2_b200.utils.tensor_utils import cumsum

    torch.manual_seed(shape[0] * shape[1])
    ip = torch.randint(high=2, size=shape).to(dtype)
    torch.testing.assert_close(cumsum(ip), torch.cumsum(ip, dim=0))


# -- test_nxd_config.py --------------------------------------------------------------------------------------------------------
def test_nxd_config_defaults():
    class Layer:
        pass

    kw = dict(tensor_parallel_size=8, pipeline_parallel_size=4, pipeline_config=None, pad_model=False, sequence_parallel=False)
    cfg = nxd.neuronx_distributed_config(optimizer_config=None, activation_checkpoint_config=Layer, model_init_config=None, **kw)
    assert cfg["optimizer_config"] == {"zero_one_enabled": False, "grad_clipping": True, "max_grad_norm": 1.0}
    assert cfg["model_init_config"] == {"sequential_move_factor": 11, "meta_device_init": False, "param_init_fn": None}
    assert cfg["pipeline_config"] is None and cfg["activation_checkpoint_config"] is Layer
    assert cfg["pad_model"] is False and cfg["sequence_parallel"] is False
    assert cfg["mixed_precision_config"] == {"use_master_weights": False, "use_fp32_grad_acc": False,
                                             "use_master_weights_in_ckpt": False}
    cfg = nxd.neuronx_distributed_config(optimizer_config={"zero_one_enabled": True}, **kw)
    assert cfg["optimizer_config"] == {"zero_one_enabled": True, "grad_clipping": True, "max_grad_norm": 1.0}
    assert cfg["mixed_precision_config"] == {"use_master_weights": True, "use_fp32_grad_acc": True,
                                             "use_master_weights_in_ckpt": False}
    init = {"meta_device_init": True, "param_init_fn": lambda m, d: None}
    cfg = nxd.neuronx_distributed_config(model_init_config=init, **kw)
    assert cfg["model_init_config"]["meta_device_init"] and cfg["model_init_config"]["param_init_fn"] is init["param_init_fn"]
    with pytest.raises(ValueError):
        nxd.neuronx_distributed_config(model_init_config={"meta_device_init": True}, **kw)
    mp = {"use_master_weights": True, "use_fp32_grad_acc": True, "use_master_weights_in_ckpt": False}
    assert nxd.neuronx_distributed_config(mixed_precision_config=dict(mp), **kw)["mixed_precision_config"] == mp


# -- test_post_partition_hooks.py ----------------------------------------------------------------------------------------------
def test_post_partition_hooks_run_once_in_order():
    from neuronx_distributed_llama3_2_b200.trainer import PostPartitionHooks

    hooks = PostPartitionHooks()
    hooks.register_post_partition_hook(lambda a, b: a * b, [5, 6])
    hooks.register_post_partition_hook(lambda a, b=1: a + b, [5], {"b": 2})
    assert len(hooks.hooks) == 2
    assert hooks.execute_all_hooks() == [30, 7]
    assert len(hooks.hooks) == 0


# -- test_get_delayed_tracing.py (disabled in the reference as well) -------------------------------------------------------------
def test_delayed_tracing_is_off():
    from neuronx_distributed_llama3_2_b200.utils.model_utils import check_delay_tracing, get_delay_tracing

    assert get_delay_tracing({"pipeline_config": {"_delay_tracing": True}}) is False
    assert check_delay_tracing({"pipeline_config": {"use_model_wrapper": True}}) is False


# -- wrapper/test_model_wrapper.py, wrapper/test_optimizer_wrapper.py ------------------------------------------------------------
def test_model_and_optimizer_wrappers():
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama
    from neuronx_distributed_llama3_2_b200.optimizer import NeuronZero1Optimizer
    from neuronx_distributed_llama3_2_b200.trainer.optimizer import NxDOptimizer

    for zero1 in (True, False):
        cfg = nxd.neuronx_distributed_config(optimizer_config={"zero_one_enabled": zero1, "grad_clipping": True,
                                                               "max_grad_norm": 1.0}, activation_checkpoint_config="full")
        model = nxd.initialize_parallel_model(cfg, lambda: LlamaForCausalLM(tiny_llama(sequence_parallel_enabled=False)),
                                              device=torch.device("cpu"))
        assert model.nxd_config == cfg and "NxDPPModel" not in str(model)
        result = model.load_state_dict(model.state_dict())
        assert result is not None and not result.missing_keys and not result.unexpected_keys
        assert all(k.startswith("model_prefix.") for k in model.state_dict(prefix="model_prefix."))
        opt = nxd.initialize_parallel_optimizer(cfg, torch.optim.AdamW, model.parameters(), lr=1e-3)
        assert opt.nxd_config == cfg and isinstance(opt, NxDOptimizer)
        assert len(list(model.parameters())) == len(opt.params) and opt.grad_norm is None
        if zero1:
            assert isinstance(opt.optimizer, NeuronZero1Optimizer)
            assert isinstance(opt.optimizer.base_optimizer, torch.optim.AdamW)
        else:
            assert isinstance(opt.optimizer, torch.optim.AdamW)
        ids = torch.randint(0, 512, (2, 16))
        model(ids, labels=ids).backward()
        opt.step()
        assert opt.grad_norm is not None and float(opt.grad_norm) > 0
        opt.zero_grad()
        sd = opt.state_dict()
        assert sd
        if zero1:       # the view reflects the live fused state: one step taken, moments non-zero
            inner = opt.optimizer.base_optimizer.state_dict()["state"]
            assert inner and all(float(st["exp_avg_sq"].abs().sum()) > 0 for st in inner.values())
