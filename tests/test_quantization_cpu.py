"""int8 weight-only quantization: observer, tensor/state-dict quantization, layer conversion and numerics
(reference ``test/unit_test/quantization/test_*.py``)."""
from __future__ import annotations

import pytest
import torch

from neuronx_distributed_llama3_2_b200 import quantization as Q
from neuronx_distributed_llama3_2_b200.parallel.layers import ColumnParallelLinear, RowParallelLinear


def test_observer_and_tensor_quantization():
    torch.manual_seed(0)
    w = torch.randn(16, 32)
    obs = Q.PerChannelAbsMaxObserver(ch_axis=0)
    obs(w[:, :16])
    obs(w)
    scale, zp = obs.calculate_qparams()
    torch.testing.assert_close(scale, w.abs().amax(1) / 127.0)
    assert int(zp.abs().sum()) == 0
    q, s = Q.quantize_per_channel_symmetric(w, 0)
    assert q.dtype == torch.int8 and s.shape == (16, 1)
    assert (Q.scale_dequantize(q, s, torch.float32) - w).abs().max() <= s.max() * 0.5 + 1e-6
    qt, st = Q.quantize_per_tensor_symmetric(w)
    assert st.shape == (1,) and (Q.scale_dequantize(qt, st, torch.float32) - w).abs().max() <= st.item() * 0.5 + 1e-6
    assert Q.direct_cast_dequantize(qt, torch.bfloat16).dtype == torch.bfloat16


def test_qint8_state_dict_adaptor():
    w = torch.randn(8, 4)
    qw = torch.quantize_per_channel(w, torch.full((8,), 0.05), torch.zeros(8, dtype=torch.int64), 0, torch.qint8)
    out = Q.convert_qint8_to_int8_state_dict({"l.weight": qw, "l.bias": torch.zeros(8)})
    assert out["l.weight"].dtype == torch.int8 and out["l.scale"].shape == (8, 1) and "l.bias" in out
    qt = torch.quantize_per_tensor(w, 0.1, 0, torch.qint8)
    out = Q.convert_qint8_to_int8_state_dict({"l.weight": qt})
    assert abs(out["l.scale"].item() - 0.1) < 1e-7


class _MLP(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.up = ColumnParallelLinear(32, 64, bias=True, gather_output=False)
        self.down = RowParallelLinear(64, 32, bias=True, input_is_parallel=True)

    def forward(self, x):
        return self.down(torch.relu(self.up(x)))


def test_convert_and_numerics_per_channel_and_per_tensor():
    torch.manual_seed(0)
    fp = _MLP()
    x = torch.randn(5, 32)
    ref = fp(x)
    for per_channel in (True, False):
        cfg = Q.get_default_per_channel_custom_qconfig_dict() if per_channel else Q.get_default_custom_qconfig_dict()
        qm = Q.convert(fp, cfg)
        assert isinstance(qm.up, Q.QuantizedColumnParallel) and isinstance(qm.down, Q.QuantizedRowParallel)
        assert isinstance(fp.up, ColumnParallelLinear)          # not in place
        sd = Q.quantize_state_dict(fp.state_dict(), per_channel=per_channel, ch_axis=0)
        assert sd["up.weight"].dtype == torch.int8 and "up.scale" in sd
        qm.load_state_dict(sd)
        out = qm(x)
        assert ((out - ref).norm() / ref.norm()).item() < (0.02 if per_channel else 0.04)


def test_state_dict_adaptor_class_reads_plain_and_packed_layouts():
    from neuronx_distributed_llama3_2_b200.quantization.quantization_layers import (
        QuantizedParallelLinearLayerStateDictAdaptor as A)

    w = torch.randn(6, 4)
    qw = torch.quantize_per_channel(w, torch.full((6,), 0.05), torch.zeros(6, dtype=torch.int64), 0, torch.qint8)
    bias = torch.arange(6.0)
    packed = {"fc._packed_params._packed_params": (qw, bias), "fc._packed_params.dtype": torch.qint8}
    assert A.get_weight_from_state_dict("fc.", packed).dtype == torch.int8
    assert torch.equal(A.get_bias_from_state_dict("fc.", packed), bias)          # the bias, not the weight
    assert A.get_scale_from_state_dict("fc.", packed).shape == (6, 1)
    plain = {}
    A.set_weight_to_state_dict("fc.", qw.int_repr(), plain)
    A.set_scale_to_state_dict("fc.", torch.full((6, 1), 0.05), plain)
    A.set_bias_to_state_dict("fc.", bias, plain)
    assert set(plain) == {"fc.weight", "fc.scale", "fc.bias"}
    assert torch.equal(A.get_weight_from_state_dict("fc.", plain), qw.int_repr())
    assert float(A.get_scale_from_state_dict("fc.", plain)[0]) == pytest.approx(0.05)
    qt = torch.quantize_per_tensor(w, 0.1, 0, torch.qint8)
    assert A.get_scale_from_state_dict("g.", {"g._packed_params._packed_params": (qt, None)}).shape == (1,)


def test_torch_ao_quantized_models_convert_to_int8_state_dicts():
    torch.manual_seed(0)
    lin = torch.nn.Sequential(torch.nn.Linear(16, 12), torch.nn.ReLU(), torch.nn.Linear(12, 4))
    for fn, scale_shape in ((Q.quantize_pytorch_model_per_channel_symmetric, (12, 1)),
                            (Q.quantize_pytorch_model_per_tensor_symmetric, (1,))):
        qm = fn(lin)
        qw = qm[0].weight()
        assert Q.extract_q_scale(qw).shape == scale_shape
        assert (qw.dequantize() - lin[0].weight).abs().max().item() < 0.01
        sd = Q.convert_qint8_to_int8_state_dict(qm.state_dict())
        assert set(sd) == {"0.weight", "0.scale", "0.bias", "2.weight", "2.scale", "2.bias"}
        assert sd["0.weight"].dtype == torch.int8 and sd["0.scale"].shape == scale_shape
        deq = sd["0.weight"].float() * sd["0.scale"]
        assert (deq - lin[0].weight).abs().max().item() < 0.01
