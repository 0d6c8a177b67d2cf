"""Throughput meter, metrics file, LR schedule and synthetic data used by the training drivers (SURVEY 5.5, 6.3;
reference ``examples/training/llama/training_utils.py:254-352``, ``tp_zero1_llama_hf_pretrain.py:61-129``, ``lr.py``)."""
from __future__ import annotations

import json
import math
import os
import queue
import time
from datetime import datetime, timezone
from typing import Any, Dict, Iterator, List, Optional

import torch


class Throughput:
    """Moving-window (default 10 steps) throughput in sequences/s =
    ``window * (batch * dp * grad_accum * logging_interval) / sum(step wall time)`` -- the reference's definition
    (``training_utils.py:329-351``) so the numbers are comparable."""

    def __init__(self, batch_size: int, world_size: int, grad_accum_usteps: int, moving_avg_window_size: int = 10,
                 logging_interval: int = 1):
        self.seqs_per_iteration = batch_size * world_size * grad_accum_usteps * logging_interval
        self.moving_avg_window_size = math.ceil(moving_avg_window_size / logging_interval)
        self.moving_avg_window: "queue.Queue[float]" = queue.Queue()
        self.window_time = 0.0
        self.start_time = time.time()

    def get_throughput(self) -> float:
        step_time = time.time() - self.start_time
        self.start_time += step_time
        self.window_time += step_time
        self.moving_avg_window.put(step_time)
        window_size = self.moving_avg_window.qsize()
        if window_size > self.moving_avg_window_size:
            self.window_time -= self.moving_avg_window.get()
            window_size -= 1
        return window_size * self.seqs_per_iteration / self.window_time


class TrainingMetrics:
    """JSON metrics file with ``MetricName / MeasuredValue / Units / Timestamp / AdditionalData`` records."""

    def __init__(self, json_file: str):
        self.json_file = json_file

    def read_modify_write_file(self, data: Any, key: str = "metrics") -> None:
        result = {}
        if os.path.exists(self.json_file):
            try:
                result = json.load(open(self.json_file))
            except json.JSONDecodeError:
                result = {}
        if key == "metrics":
            result.setdefault("metrics", [])
            names = {m["MetricName"] for m in data}
            result["metrics"] = [m for m in result["metrics"] if m["MetricName"] not in names] + list(data)
        else:
            result[key] = data
        with open(self.json_file, "w") as f:
            json.dump(result, f, indent=1)

    def store_metrics(self, metrics: List[Dict[str, Any]]) -> None:
        self.read_modify_write_file(metrics, "metrics")

    def store_parameters(self, parameters: Dict[str, Any]) -> None:
        self.read_modify_write_file(parameters, "parameters")

    @staticmethod
    def metric(name: str, value: float, units: str, additional_data: Optional[Dict[str, Any]] = None) -> Dict[str, Any]:
        return {"MetricName": name, "MeasuredValue": value, "Units": units,
                "Timestamp": datetime.now(timezone.utc).isoformat(), "AdditionalData": additional_data or {}}


class CosineAnnealing(torch.optim.lr_scheduler.LRScheduler):
    """Linear warm-up then cosine decay to ``min_lr``, optional constant tail (reference ``lr.py``)."""

    def __init__(self, optimizer, max_steps: int, warmup_steps: int = 0, constant_steps: int = 0, min_lr: float = 0.0,
                 last_epoch: int = -1):
        self.max_steps, self.warmup_steps, self.constant_steps, self.min_lr = max_steps, warmup_steps, constant_steps, min_lr
        self.decay_steps = max(max_steps - constant_steps - warmup_steps, 1)
        super().__init__(optimizer, last_epoch)

    def get_lr(self):
        step = self.last_epoch
        out = []
        for base in self.base_lrs:
            if self.warmup_steps > 0 and step <= self.warmup_steps:
                out.append(base * (step + 1) / (self.warmup_steps + 1))
            elif step > self.warmup_steps + self.decay_steps:
                out.append(self.min_lr)
            else:
                t = (step - self.warmup_steps) / self.decay_steps
                out.append(self.min_lr + 0.5 * (base - self.min_lr) * (1 + math.cos(math.pi * t)))
        return out


class SyntheticTokens:
    """Endless synthetic ``input_ids`` batches in pinned host memory with a side-stream prefetch to the device (the
    B200 stand-in for ``MpDeviceLoader``)."""

    def __init__(self, vocab_size: int, batch: int, seq_len: int, device, seed: int = 0, n_batches: int = 8):
        g = torch.Generator().manual_seed(seed)
        self.host = [torch.randint(0, vocab_size, (batch, seq_len), generator=g) for _ in range(n_batches)]
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        if self.cuda:
            self.host = [h.pin_memory() for h in self.host]
            self.stream = torch.cuda.Stream(device=self.device)
        self.i = 0
        self._next = None
        self._prefetch()

    def _prefetch(self) -> None:
        h = self.host[self.i % len(self.host)]
        self.i += 1
        if self.cuda:
            with torch.cuda.stream(self.stream):
                self._next = h.to(self.device, non_blocking=True)
        else:
            self._next = h

    def __iter__(self) -> Iterator[torch.Tensor]:
        return self

    def __next__(self) -> torch.Tensor:
        if self.cuda:
            torch.cuda.current_stream(self.device).wait_stream(self.stream)
        out = self._next
        self._prefetch()
        return out


# -- the remaining helpers of the reference's example utilities (``examples/training/llama/training_utils.py``, ``lr.py``) ----
class WarmupAnnealHoldPolicy(torch.optim.lr_scheduler.LRScheduler):
    """Base schedule: linear warm-up for ``warmup_steps`` (or ``warmup_ratio * max_steps``), a decay phase implemented by
    ``_get_lr(step)`` in subclasses, then ``min_lr`` held for ``constant_steps`` (reference ``lr.py:25-112``)."""

    def __init__(self, optimizer, *, max_steps: int, warmup_steps: Optional[int] = None, warmup_ratio: Optional[float] = None,
                 constant_steps: Optional[int] = None, constant_ratio: Optional[float] = None, min_lr: float = 0.0,
                 last_epoch: int = -1):
        assert not (warmup_steps is not None and warmup_ratio is not None), "give warmup_steps OR warmup_ratio"
        assert not (constant_steps is not None and constant_ratio is not None), "give constant_steps OR constant_ratio"
        self.max_steps, self.min_lr = max_steps, min_lr
        self.warmup_steps = warmup_steps if warmup_steps is not None else int((warmup_ratio or 0.0) * max_steps)
        self.constant_steps = constant_steps if constant_steps is not None else int((constant_ratio or 0.0) * max_steps)
        self.decay_steps = max_steps - (self.constant_steps + self.warmup_steps)
        super().__init__(optimizer, last_epoch)

    def get_lr(self):
        step = self.last_epoch
        if self.warmup_steps > 0 and step <= self.warmup_steps:
            return [base * (step + 1) / (self.warmup_steps + 1) for base in self.base_lrs]
        if self.constant_steps > 0 and self.warmup_steps + self.decay_steps < step <= self.max_steps:
            return [self.min_lr for _ in self.base_lrs]
        if step > self.max_steps:
            return [self.min_lr for _ in self.base_lrs]
        return self._get_lr(step)

    def _get_lr(self, step: int):
        return list(self.base_lrs)


def get_learning_rate_scheduler(optimizer, args, last_epoch: int = -1) -> CosineAnnealing:
    """``CosineAnnealing`` from ``args.max_steps / min_lr / warmup_steps / constant_steps`` (reference ``:65-74``)."""
    return CosineAnnealing(optimizer, max_steps=args.max_steps, min_lr=getattr(args, "min_lr", 0.0),
                           warmup_steps=getattr(args, "warmup_steps", 0),
                           constant_steps=getattr(args, "constant_steps", 0), last_epoch=last_epoch)


def get_param_groups_by_weight_decay(model, weight_decay: float = 0.01) -> List[Dict[str, Any]]:
    """Two groups: weights with decay; biases and norm scales without (reference ``:77-96``).  A partitioned pipeline
    model contributes its LOCAL parameters only."""
    if getattr(model, "partitioned", False) and hasattr(model, "local_named_parameters"):
        named = list(model.local_named_parameters())
    else:
        named = list(model.named_parameters())
    no_decay = ("bias", "LayerNorm", "layernorm", "layer_norm", "norm.weight")

    def is_plain(n, p):          # by name as the reference, and every 1-D tensor (norm scales, biases) whatever its name
        return p.dim() <= 1 or any(k in n for k in no_decay)

    decay = [p for n, p in named if p.requires_grad and not is_plain(n, p)]
    plain = [p for n, p in named if p.requires_grad and is_plain(n, p)]
    return [{"params": decay, "weight_decay": weight_decay}, {"params": plain, "weight_decay": 0.0}]


def create_partition(num_hidden_layers: int, pipeline_parallel_size: int, prefix: str = "model.layers.") -> List[str]:
    """Names of the layers AFTER which the pipeline is cut: even split, remainder to the later stages (reference
    ``:210-221``)."""
    from ..pipeline.partition import create_partitions

    return [f"{prefix}{i}" for i in create_partitions(pipeline_parallel_size, num_hidden_layers)]


def get_sin_cos_matrix(config):
    """``(sin, cos)`` rotary tables ``[max_position_embeddings, head_dim / 2]`` in fp32 (reference ``:224-232``)."""
    head_dim = getattr(config, "head_dim", None) or config.hidden_size // config.num_attention_heads
    base = getattr(config, "rope_theta", 10000.0)
    inv_freq = 1.0 / (base ** (torch.arange(0, head_dim, 2).float() / head_dim))
    t = torch.arange(config.max_position_embeddings, dtype=torch.float32)
    freqs = torch.outer(t, inv_freq)
    return freqs.sin(), freqs.cos()


def get_dtype(model) -> str:
    """``torch.float32`` / ``torch.bfloat16`` ... of the model's parameters; with master weights kept by the optimizer
    the model dtype is what the forward runs in (reference ``:235-246``)."""
    cfg = getattr(model, "nxd_config", None) or {}
    if cfg.get("mixed_precision_config", {}).get("use_master_weights") and torch.cuda.is_available():
        return "torch.bfloat16"
    p = next(iter(model.parameters()), None)
    return str(p.dtype) if p is not None else "torch.float32"


def print_logs(loss, global_norm, args, throughput, logger, total_steps: int, current_lr: float,
               input_ids: torch.Tensor, start: float) -> None:
    """One training log line + TensorBoard scalars (reference ``:249-252``).  ``throughput`` is the seq/s number or a
    ``Throughput`` meter, ``logger`` a ``TrainingLogger``."""
    tps = throughput.get_throughput() if hasattr(throughput, "get_throughput") else float(throughput)
    gn = float(global_norm) if global_norm is not None else None
    logger.log(0, total_steps, float(loss), current_lr, tps, gn)


def load_events(event_file: str) -> Dict[str, Any]:
    """Scalars of a TensorBoard event file, ``{tag: [ScalarEvent, ...]}`` (reference ``tp_zero1_llama_hf_pretrain/logger.py:11-19``)."""
    from tensorboard.backend.event_processing.event_accumulator import EventAccumulator

    acc = EventAccumulator(event_file)
    acc.Reload()
    return {tag: acc.Scalars(tag) for tag in acc.Tags()["scalars"]}


class TrainingLogger:
    """Console line + TensorBoard scalars ``step loss / learning rate / throughput / grad-norm`` under a run directory
    named after the hyper-parameters (reference ``tp_zero1_llama_hf_pretrain/logger.py:22-75``).  Create it on the rank
    that should log (global rank 0, or the last PP stage's dp0/tp0 rank)."""

    def __init__(self, args, world_size: int, model_dtype: str):
        self.throughputs: List[float] = []
        run = (f"nxd_tblogs_{time.strftime('%m%d%y_%H%M')}_{str(model_dtype).replace('torch.', '')}_w{world_size}"
               f"_lr{getattr(args, 'lr', 0)}_bs{getattr(args, 'batch_size', 0)}_acc{getattr(args, 'grad_accum_usteps', 1)}"
               f"_warmup{getattr(args, 'warmup_steps', 0)}_max{getattr(args, 'max_steps', 0)}_{self.get_instance_type()}")
        self.log_dir = os.path.join(getattr(args, "output_dir", "."), run)
        try:
            from torch.utils.tensorboard import SummaryWriter

            self.tb = SummaryWriter(self.log_dir)
        except Exception:  # noqa: BLE001 - tensorboard is optional
            self.tb = None

    @staticmethod
    def get_instance_type() -> str:
        if torch.cuda.is_available():
            return torch.cuda.get_device_name(0).replace(" ", "_")
        return os.environ.get("HOSTNAME", "cpu")

    def log(self, epoch: int, step: int, step_loss: float, learning_rate: float, throughput: float,
            grad_norm: Optional[float] = None) -> None:
        gn = f"grad-norm : {grad_norm:.4f}" if grad_norm is not None else ""
        print(f"LOG {time.asctime()} - ({epoch}, {step}) step_loss : {step_loss:.4f} learning_rate : {learning_rate:.2e} "
              f"throughput : {throughput:.2f} seq/s {gn}", flush=True)
        if self.tb is not None:
            self.tb.add_scalar("step loss", step_loss, step)
            self.tb.add_scalar("learning rate", learning_rate, step)
            self.tb.add_scalar("throughput", throughput, step)
            if grad_norm is not None:
                self.tb.add_scalar("grad-norm", grad_norm, step)
        self.throughputs.append(throughput)

    def close(self) -> None:
        if self.tb is not None:
            self.tb.flush()
            self.tb.close()


Logger = TrainingLogger  # the reference's example class name


def get_mixed_precision_config(use_gpu_compatible_precision) -> Dict[str, bool]:
    """fp32 master weights + fp32 gradient accumulation when GPU-compatible precision is requested (reference ``:354-359``)."""
    on = bool(use_gpu_compatible_precision)
    return {"use_master_weights": on, "use_fp32_grad_acc": on, "use_master_weights_in_ckpt": False}
