"""A minimal fit loop that drives the Lightning-protocol objects of this package when Lightning itself is absent."""
from __future__ import annotations

from typing import Iterable, Optional


class NxDLightningLoop:
    def __init__(self, strategy, logger=None, progress_bar=None, max_steps: int = 10, checkpoint_dir: Optional[str] = None,
                 checkpoint_every: int = 0):
        self.strategy, self.logger, self.progress_bar = strategy, logger, progress_bar
        self.max_steps, self.checkpoint_dir, self.checkpoint_every = max_steps, checkpoint_dir, checkpoint_every

    def fit(self, module, dataloader: Iterable) -> None:
        self.strategy.setup_distributed()
        module.trainer = self
        module.setup("fit")
        module.configure_optimizers()
        if self.progress_bar is not None:
            if hasattr(self.progress_bar, "setup"):
                self.progress_bar.setup(self, module, "fit")
            self.progress_bar.on_train_start(total=self.max_steps)
        it = iter(dataloader)
        while module.global_step_ < self.max_steps:
            batch = self.strategy.batch_to_device(next(it))
            loss = module.training_step(batch)
            if module._micro == 0:
                if self.logger is not None and loss is not None and module.should_print():
                    self.logger.log_metrics({"step loss": float(loss)}, step=module.global_step_)
                if self.progress_bar is not None:
                    self.progress_bar.on_train_batch_end()
                if self.checkpoint_dir and self.checkpoint_every and module.global_step_ % self.checkpoint_every == 0:
                    self.strategy.save_checkpoint({"model": module.state_dict()},
                                                  f"{self.checkpoint_dir}/step_{module.global_step_}")
        if self.progress_bar is not None:
            self.progress_bar.on_train_end()
        if self.logger is not None:
            self.logger.finalize()
