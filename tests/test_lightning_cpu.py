"""Lightning-protocol integration driven by the local fit loop (Lightning is not installed in this image)."""
from __future__ import annotations

import os
import tempfile

import torch


def test_lightning_module_strategy_loop_single_process():
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.lightning import (NeuronCheckpointIO, NeuronLTModule, NeuronTensorBoardLogger,
                                                             NeuronTQDMProgressBar, NeuronXLAPrecisionPlugin,
                                                             NeuronXLAStrategy, NxDLightningLoop)
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama
    from neuronx_distributed_llama3_2_b200.utils.training_utils import CosineAnnealing, SyntheticTokens

    cfg = nxd.neuronx_distributed_config(optimizer_config={"zero_one_enabled": True, "grad_clipping": True,
                                                           "max_grad_norm": 1.0})
    torch.manual_seed(0)
    module = NeuronLTModule(cfg, torch.optim.AdamW, CosineAnnealing, model_fn=lambda **kw: LlamaForCausalLM(tiny_llama()),
                            opt_kwargs={"lr": 5e-3}, scheduler_kwargs={"max_steps": 20, "warmup_steps": 2},
                            grad_accum_steps=2, device=torch.device("cpu"))
    with tempfile.TemporaryDirectory() as d:
        strategy = NeuronXLAStrategy(nxd_config=cfg, save_load_xser=False)
        assert isinstance(strategy.checkpoint_io, NeuronCheckpointIO)
        logger = NeuronTensorBoardLogger(d)
        loop = NxDLightningLoop(strategy, logger=logger, progress_bar=NeuronTQDMProgressBar(), max_steps=4,
                                checkpoint_dir=os.path.join(d, "ck"), checkpoint_every=2)
        data = SyntheticTokens(512, 2, 32, "cpu", n_batches=1)
        loop.fit(module, data)
        assert module.global_step_ == 4 and module._logged["loss"] < 7.0 and "grad_norm" in module._logged
        assert os.path.exists(os.path.join(d, "ck", "step_2", "tp_rank_00_pp_rank_00", "checkpoint.pt"))
        assert os.path.isdir(logger.log_dir)
        loaded = strategy.load_checkpoint(os.path.join(d, "ck", "step_4"))
        assert "model" in loaded and "model.norm.weight" in loaded["model"]
    with NeuronXLAPrecisionPlugin(mode="autocast").forward_context():
        pass
    assert strategy.distributed_sampler_kwargs == {"num_replicas": 1, "rank": 0}
    from neuronx_distributed_llama3_2_b200.lightning import NeuronXLAAccelerator
    reg = {}
    NeuronXLAAccelerator.register_accelerators(reg)
    assert reg["neuron"] is NeuronXLAAccelerator and reg["nxd_cuda"] is NeuronXLAAccelerator
    bar = NeuronTQDMProgressBar()
    bar.setup(loop, module, "fit")
    assert bar.is_enabled            # dp 0 / tp 0 / last pp stage of a 1-process job
    bar.disable()
    assert not bar.is_enabled
