"""Latency collectors and report (reference ``examples/inference/modules/benchmark.py:9-72``): p50 / p90 /
p95 / p99 / p100 / average latency in ms and throughput = ``n_runs * max_length * batch / sum(latency)``.  Latencies are
taken with CUDA events on the launching stream (the reference uses ``time.time()`` around forward hooks)."""
from __future__ import annotations

import time
from typing import Dict, List, Optional

import numpy as np
import torch

BENCHMARK_REPORT_FILENAME = "benchmark_report.json"


class LatencyCollector:
    """Collects one latency per ``pre_hook`` / ``hook`` pair.  On CUDA the pair records events; ``finish()`` resolves
    them to milliseconds after a single synchronisation (no per-call host sync in the measured loop)."""

    def __init__(self, cuda: Optional[bool] = None):
        self.cuda = torch.cuda.is_available() if cuda is None else cuda
        self.latency_list: List[float] = []
        self._events = []
        self._t0 = None

    def pre_hook(self, *_):
        if self.cuda:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self._t0 = e
        else:
            self._t0 = time.perf_counter()

    def hook(self, *_):
        if self.cuda:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self._events.append((self._t0, e))
        else:
            self.latency_list.append((time.perf_counter() - self._t0) * 1e3)

    def finish(self) -> None:
        if self.cuda and self._events:
            torch.cuda.synchronize()
            self.latency_list += [a.elapsed_time(b) for a, b in self._events]
            self._events = []

    def percentile(self, p: float) -> float:
        self.finish()
        return float(np.percentile(np.asarray(self.latency_list), p)) if self.latency_list else float("nan")


def generate_report(latency_list_ms: List[float], max_length: int, max_batch_size: int, n_runs: int = 20) -> Dict:
    lat = np.asarray(latency_list_ms, dtype=np.float64)
    total_s = float(lat.sum()) / 1e3
    rep = {f"latency_ms_p{p}": float(np.percentile(lat, p)) for p in (50, 90, 95, 99, 100)}
    rep["latency_ms_avg"] = float(lat.mean())
    rep["throughput"] = (n_runs * max_length * max_batch_size) / total_s if total_s > 0 else float("nan")
    return rep


class Benchmark:
    def __init__(self, benchmark_func, input_param=None, config=None, num_runs: int = 20, preprocess_func=None):
        self.benchmark_func, self.input_param, self.config = benchmark_func, input_param, config
        self.num_runs, self.preprocess_func = num_runs, preprocess_func
        self.latency_list: List[float] = []

    def run(self) -> Dict:
        c = LatencyCollector()
        for _ in range(self.num_runs):
            if self.preprocess_func:
                self.preprocess_func()
            c.pre_hook()
            if isinstance(self.input_param, (tuple, list)):
                self.benchmark_func(*self.input_param)
            elif isinstance(self.input_param, dict):
                self.benchmark_func(**self.input_param)
            else:
                self.benchmark_func()
            c.hook()
        c.finish()
        self.latency_list = c.latency_list
        return generate_report(c.latency_list, getattr(self.config, "seq_len", 1),
                               getattr(self.config, "max_batch_size", 1), self.num_runs)
