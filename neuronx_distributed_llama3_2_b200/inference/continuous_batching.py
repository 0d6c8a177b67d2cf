"""Continuous batching over the paged KV pool: requests of different lengths join and leave the running batch between
decode steps (the reference's ``is_continuous_batching`` mode passes ``seq_ids`` so that a step's rows map to arbitrary
cache lines, ``examples/inference/modules/model_wrapper.py``; it leaves admission to the caller -- here a native
scheduler decides it from the free pages).

One iteration = (1) admit waiting requests while a cache line and pages are free and run their prompts (context
encoding, one request per call, on the line the scheduler assigned); (2) one token-generation call for every request
that was already running, rows = requests, ``seq_ids`` = their lines; (3) retire finished requests (pages return to the
pool), grow the others' reservations, preempt the youngest when the pool runs dry (they re-run their prompt later).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

from ..runtime.native import NativeScheduler


@dataclass
class _Request:
    rid: int
    prompt: torch.Tensor                     # [T] token ids
    max_new_tokens: int
    line: int = -1
    output: List[int] = field(default_factory=list)


class ContinuousBatcher:
    """``model`` is a KV-cache inference model (``KVCacheInferenceMixin``) whose ``InferenceConfig`` has
    ``paged_kv=True``, ``num_kv_pages`` set and ``is_continuous_batching=True``.  ``pad_decode_to`` pads the decode batch
    to a fixed size by repeating row 0 (identical rows write identical K/V, so the duplicate is harmless) -- what a
    captured token-generation graph of that batch size needs."""

    def __init__(self, model, max_running: Optional[int] = None, eos_token_id: Optional[int] = None,
                 pad_decode_to: Optional[int] = None, context_buckets: Optional[List[int]] = None):
        self.model = model
        if model.kv_cache is None:
            model.init_kv_cache()
        self.cache = model.kv_cache
        assert self.cache.pool is not None, "continuous batching needs InferenceConfig(paged_kv=True, num_kv_pages=N)"
        ic = model.inference_config
        self.max_running = max_running or self.cache.lines
        self.sched = NativeScheduler(self.cache.pool, self.max_running)
        self.eos = eos_token_id if eos_token_id is not None else ic.eos_token_id
        self.pad_decode_to = pad_decode_to
        self.buckets = sorted(context_buckets) if context_buckets else None
        self.pad_id = ic.pad_token_id
        self.requests: Dict[int, _Request] = {}
        self.running: List[int] = []                      # request ids in admission order
        self._next_id = 0
        self.stats = {"iterations": 0, "prefills": 0, "decode_calls": 0, "decode_rows": 0, "preemptions": 0,
                      "max_running": 0}

    def submit(self, prompt: torch.Tensor, max_new_tokens: int) -> int:
        rid = self._next_id
        self._next_id += 1
        prompt = prompt.reshape(-1)
        self.sched.submit(rid, int(prompt.numel()), int(max_new_tokens))
        self.requests[rid] = _Request(rid, prompt, int(max_new_tokens))
        return rid

    # -- one model call each --------------------------------------------------------------------------------------
    def _prefill(self, req: _Request) -> int:
        dev = self.cache.block_table.device
        T = int(req.prompt.numel())
        Tb = next((b for b in self.buckets if b >= T), T) if self.buckets else T
        ids = torch.full((1, Tb), self.pad_id, dtype=torch.long, device=dev)
        ids[0, :T] = req.prompt.to(dev)
        if Tb > T:                                            # the padded tail is written to the cache too
            self.cache.reserve(req.line, Tb + 1)
        pos = torch.arange(Tb, device=dev).unsqueeze(0)
        tok = self.model.forward_infer(ids, pos, seq_ids=torch.tensor([req.line], dtype=torch.int32, device=dev),
                                       last_token_index=torch.tensor([T - 1], device=dev))
        self.stats["prefills"] += 1
        return int(tok[0])

    def _decode(self, reqs: List[_Request]) -> List[int]:
        dev = self.cache.block_table.device
        rows = list(reqs)
        n = len(rows)
        if self.pad_decode_to and n < self.pad_decode_to:
            rows = rows + [rows[0]] * (self.pad_decode_to - n)
        toks = torch.tensor([[r.output[-1]] for r in rows], dtype=torch.long, device=dev)
        pos = torch.tensor([[int(r.prompt.numel()) + len(r.output) - 1] for r in rows], dtype=torch.long, device=dev)
        lines = torch.tensor([r.line for r in rows], dtype=torch.int32, device=dev)
        out = self.model.forward_infer(toks, pos, seq_ids=lines)
        self.stats["decode_calls"] += 1
        self.stats["decode_rows"] += n
        return [int(t) for t in out[:n]]

    # -- the loop ------------------------------------------------------------------------------------------------------
    def step(self) -> List[int]:
        """One iteration; returns the ids of the requests that finished in it."""
        self.stats["iterations"] += 1
        before = list(self.running)
        for rid, line in self.sched.admit():
            req = self.requests[rid]
            req.line, req.output = line, []
            self.cache.sync_lines([line])
            req.output.append(self._prefill(req))
            self.running.append(rid)
        if before:
            for req, tok in zip([self.requests[r] for r in before], self._decode([self.requests[r] for r in before])):
                req.output.append(tok)
        self.stats["max_running"] = max(self.stats["max_running"], len(self.running))
        hit_eos = [rid for rid in self.running if self.eos is not None and self.requests[rid].output[-1] == self.eos]
        retired, preempted, _ = self.sched.step(hit_eos)
        for rid in preempted:                               # its pages are gone: it starts over when re-admitted
            self.requests[rid].output = []
            self.requests[rid].line = -1
            self.stats["preemptions"] += 1
        gone = set(retired) | set(preempted)
        self.running = [r for r in self.running if r not in gone]
        self.cache.sync_lines([self.requests[r].line for r in self.running])      # reservations may have grown
        return retired

    def run(self, max_iterations: int = 1_000_000) -> Dict[int, List[int]]:
        """Serve everything submitted so far; ``{request id: generated token ids}``."""
        it = 0
        while (self.running or self.sched.num_waiting) and it < max_iterations:
            self.step()
            it += 1
        return {rid: list(r.output) for rid, r in self.requests.items()}
