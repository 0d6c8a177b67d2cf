"""Chrome-trace timeline of pipeline tasks (reference ``src/neuronx_distributed/utils/timeline.py:14-137``
and ``pipeline/timeline.py:10-21``).  Events are ``ph:"B"/"E"`` pairs with ``pid`` = pp rank; on CUDA an NVTX range is
opened as well so the same labels show up in Nsight."""
from __future__ import annotations

import json
import os
import time
from typing import List, Optional

import torch


class Timeline:
    def __init__(self, trace_file_path: Optional[str], rank: int = 0):
        self.path = trace_file_path
        self.rank = rank
        self.enabled = trace_file_path is not None
        self.events: List[dict] = []
        self._nvtx = torch.cuda.is_available()
        self.step = 0

    @property
    def should_record(self) -> bool:
        """Whether this rank records events (reference ``utils/timeline.py:39-41``; subclasses narrow it)."""
        return self.enabled

    def mark_event_start(self, label: str) -> None:
        if not self.should_record:
            return
        if self._nvtx:
            torch.cuda.nvtx.range_push(label)
        self.events.append({"name": label, "ph": "B", "ts": time.time() * 1e6, "pid": self.rank, "tid": 0})

    def mark_event_end(self, label: str) -> None:
        if not self.should_record:
            return
        if self._nvtx:
            torch.cuda.nvtx.range_pop()
        self.events.append({"name": label, "ph": "E", "ts": time.time() * 1e6, "pid": self.rank, "tid": 0})

    def mark_step_end(self, gather_group=None) -> None:
        """Gather the step's events from all PP ranks (object all-gather) and append them to the trace file."""
        if not self.should_record:
            return
        import torch.distributed as dist

        events = self.events
        self.events = []
        if dist.is_initialized() and gather_group is not None:
            gathered = [None] * dist.get_world_size(group=gather_group)
            dist.all_gather_object(gathered, events, group=gather_group)
            if dist.get_rank(group=gather_group) != 0:
                return
            events = [e for part in gathered for e in part]
        prior = []
        if os.path.exists(self.path):
            try:
                prior = json.load(open(self.path))
            except Exception:  # noqa: BLE001
                prior = []
        with open(self.path, "w") as f:
            json.dump(prior + events, f)
        self.step += 1


class PPTimeline(Timeline):
    """Records only on dp rank 0 / tp rank 0 (reference ``pipeline/timeline.py:10-21``)."""

    def __init__(self, trace_file_path, pp_rank: int):
        from ..parallel import parallel_state as ps

        on = trace_file_path is not None
        if on and ps.model_parallel_is_initialized():
            on = ps.get_data_parallel_rank() == 0 and ps.get_tensor_model_parallel_rank() == 0
        super().__init__(trace_file_path if on else None, pp_rank)

    @property
    def should_record(self) -> bool:
        """Enabled and on the (dp 0, tp 0) replica: one track per pipeline stage (reference ``pipeline/timeline.py:16-18``)."""
        return self.enabled
