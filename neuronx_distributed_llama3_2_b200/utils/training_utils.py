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
