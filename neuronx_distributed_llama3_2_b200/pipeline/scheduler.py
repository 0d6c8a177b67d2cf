"""Pipeline schedules: sequences of tasks a stage executes for one batch.

Reference: ``src/neuronx_distributed/pipeline/scheduler.py:4-685`` -- same task vocabulary and the same
externally visible ordering rules:

* a *Preprocess* task receives (fwd: activations from the previous stage, bwd: grads from the next stage),
  a *Step* task computes, a *Postprocess* task sends;
* 1F1B: ``pp - rank - 1`` warm-up forwards, then one-forward-one-backward, then the remaining backwards; in the
  steady state a stage **receives the backward grad before it sends the forward activation of the previous
  micro-batch** (reference ``:226-233``), which keeps a blocking two-party exchange deadlock free;
* the last step of a training schedule is ``ReduceGradsTask``.

The schedules are pure Python (no torch) and are generated from an explicit per-stage operation list
(``_op_sequence``) instead of the reference's step-index arithmetic; ``tests/test_pipeline_scheduler.py`` replays all
stages of a schedule against each other to prove that every send meets its receive in order.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Iterator, List, Tuple


class PipelineTask:
    def __init__(self, mb: int, model_chunk: int = 0, graph_break: bool = True):
        self.mb = mb
        self.model_chunk = model_chunk
        self.graph_break = graph_break

    def __eq__(self, other) -> bool:
        return (type(self) is type(other) and self.mb == other.mb and self.model_chunk == other.model_chunk
                and self.graph_break == other.graph_break)

    def __hash__(self):
        return hash((type(self).__name__, self.mb, self.model_chunk, self.graph_break))

    def __repr__(self) -> str:
        return (f"{type(self).__name__}_microbatch_{self.mb}_modelchunk_{self.model_chunk}"
                f"_graphbreak_{self.graph_break}")


class ForwardStepTask(PipelineTask):
    pass


class ForwardPreprocessTask(PipelineTask):
    pass


class ForwardPostprocessTask(PipelineTask):
    pass


class BackwardStepTask(PipelineTask):
    pass


class BackwardPreprocessTask(PipelineTask):
    pass


class BackwardPostprocessTask(PipelineTask):
    pass


class PostProcessTask:
    """Runs once after the pipeline drained."""

    def __init__(self, graph_break: bool = True):
        self.mb = -1
        self.model_chunk = -1
        self.graph_break = graph_break

    def __eq__(self, other) -> bool:
        return type(self) is type(other)

    def __hash__(self):
        return hash(type(self).__name__)


class ReduceGradsTask(PostProcessTask):
    def __repr__(self) -> str:
        return "ReduceGradsTask"


class PipeSchedule(ABC):
    """Generator of per-step task lists for stage ``stage_id`` of ``stages``."""

    def __init__(self, num_microbatches: int, stages: int, stage_id: int):
        self.num_microbatches = num_microbatches
        self.stages = stages
        self.stage_id = stage_id
        self.prev_stage = stage_id - 1
        self.next_stage = stage_id + 1

    @abstractmethod
    def steps(self) -> Iterator[List[object]]:
        ...

    def _valid_micro_batch(self, mb: int) -> bool:
        return 0 <= mb < self.num_microbatches

    def _valid_stage(self, stage: int) -> bool:
        return 0 <= stage < self.stages

    @property
    def stage(self) -> int:
        return self.stage_id

    @property
    def num_stages(self) -> int:
        return self.stages

    @property
    def is_first_stage(self) -> bool:
        return self.stage_id == 0

    @property
    def is_last_stage(self) -> bool:
        return self.stage_id == self.stages - 1

    def __iter__(self):
        self.it = None
        return self

    def __next__(self):
        if self.it is None:
            self.it = self.steps()
        return next(self.it)


class InferenceSchedule(PipeSchedule):
    """Forward only: every micro-batch is receive -> compute -> send.  With virtual stages the model chunks are
    walked one after the other (all micro-batches through chunk 0, then chunk 1, ...)."""

    def __init__(self, num_microbatches: int, stages: int, stage_id: int, num_model_chunks: int = 1):
        super().__init__(num_microbatches, stages, stage_id)
        self.num_model_chunks = num_model_chunks

    def steps(self):
        for chunk in range(self.num_model_chunks):
            for mb in range(self.num_microbatches):
                yield [ForwardPreprocessTask(mb, chunk), ForwardStepTask(mb, chunk), ForwardPostprocessTask(mb, chunk)]


class Train1F1BSchedule(PipeSchedule):
    """Non-interleaved 1F1B.  Example, 4 stages x 6 micro-batches::

        stage0  F0 F1 F2 F3 B0 F4 B1 F5 B2 B3 B4 B5
        stage1     F0 F1 F2 B0 F3 B1 F4 B2 F5 B3 B4 B5
        stage2        F0 F1 B0 F2 B1 F3 B2 F4 B3 F5 B4 B5
        stage3           F0 B0 F1 B1 F2 B2 F3 B3 F4 B4 F5 B5
    """

    def __init__(self, num_microbatches: int, stages: int, stage_id: int):
        super().__init__(num_microbatches, stages, stage_id)
        self.get_microbatche_schedule()

    def get_microbatche_schedule(self) -> None:
        """Phase sizes of this stage: warm-up forwards, steady 1F1B pairs, cool-down backwards (method name, typo
        included, from the reference ``scheduler.py:179-184``)."""
        self.num_warmup_steps = min(self.stages - self.stage_id - 1, self.num_microbatches)
        self.num_steady_state_microbatches = self.num_microbatches - self.num_warmup_steps
        self.num_remaining_microbatches = self.num_warmup_steps

    def _op_sequence(self) -> List[Tuple[str, int]]:
        """[('F'|'B', micro_batch)] in execution order."""
        w, n = self.num_warmup_steps, self.num_microbatches
        ops: List[Tuple[str, int]] = [("F", i) for i in range(w)]
        for k in range(self.num_steady_state_microbatches):
            ops.append(("F", w + k))
            ops.append(("B", k))
        ops += [("B", k) for k in range(self.num_steady_state_microbatches, n)]
        return ops

    def steps(self):
        has_prev, has_next = self._valid_stage(self.prev_stage), self._valid_stage(self.next_stage)
        w = self.num_warmup_steps
        pending_send = None  # forward output of the steady phase, sent together with the next backward's receive
        for kind, mb in self._op_sequence():
            cmds: List[object] = []
            if kind == "F":
                cmds.append(ForwardPreprocessTask(mb))
                cmds.append(ForwardStepTask(mb))
                if has_next:
                    if mb < w:
                        cmds.append(ForwardPostprocessTask(mb))  # warm-up: send right away
                    else:
                        pending_send = mb                        # steady: after the next recv-bwd
            else:
                if has_next:
                    cmds.append(BackwardPreprocessTask(mb))
                    if pending_send is not None:
                        cmds.append(ForwardPostprocessTask(pending_send))
                        pending_send = None
                cmds.append(BackwardStepTask(mb))
                if has_prev:
                    cmds.append(BackwardPostprocessTask(mb))
            yield cmds
        yield [ReduceGradsTask()]


class TrainInterleavedSchedule(PipeSchedule):
    """Interleaved (virtual-pipeline) 1F1B after Megatron/Apex (reference ``scheduler.py:256-541``): every rank owns
    ``num_model_chunks`` model chunks; micro-batches advance through chunk 0 of all ranks, then chunk 1, ...
    ``num_microbatches`` must be a multiple of the pipeline size.

    Each yielded step contains the compute task plus the receives/sends adjacent to it.  ``fused_send_recv`` and
    ``fused_fwd_bwd`` only influence ``graph_break`` flags (kept for API parity: with CUDA streams there is no
    graph cut, the flags mark where the engine may batch p2p ops)."""

    def __init__(self, num_microbatches: int, num_model_chunks: int, stages: int, stage_id: int,
                 deallocate_send_tensors: bool = True, fused_send_recv: bool = False, fused_fwd_bwd: bool = False,
                 use_odd_even_scheduler: bool = False):
        super().__init__(num_microbatches, stages, stage_id)
        if num_microbatches % stages != 0:
            raise ValueError("the interleaved schedule needs num_microbatches to be a multiple of the pipeline size")
        self.num_model_chunks = num_model_chunks
        self.deallocate_send_tensors = deallocate_send_tensors
        self.fused_send_recv = fused_send_recv
        self.fused_fwd_bwd = fused_fwd_bwd
        self.use_odd_even_scheduler = use_odd_even_scheduler
        self.get_step_schedule()

    def get_step_schedule(self) -> None:
        """Warm-up / steady / cool-down sizes in virtual micro-batches (reference ``scheduler.py:296-317``): all forwards
        first when there are exactly as many micro-batches as stages, otherwise ``2 (stages - stage - 1)`` plus one round
        per extra model chunk."""
        total = self.num_microbatches * self.num_model_chunks
        self.total_num_microbatches = self.num_microbatches_steps = total
        if self.num_microbatches == self.stages:
            self.num_warmup_microbatches = total
        else:
            self.num_warmup_microbatches = min((self.stages - self.stage_id - 1) * 2
                                               + (self.num_model_chunks - 1) * self.stages, total)
        self.num_microbatches_remaining = total - self.num_warmup_microbatches
        self.num_warmup_steps = self.num_warmup_microbatches
        self.num_steady_state_steps = self.num_microbatches_remaining
        self.num_remaining_steps = self.num_warmup_microbatches

    # -- index helpers (virtual micro-batch k -> (chunk, micro-batch)) ------------------------------------------
    def get_model_chunk_id(self, k: int, forward: bool) -> int:
        group = k % (self.stages * self.num_model_chunks)
        chunk = group // self.stages
        return chunk if forward else self.num_model_chunks - 1 - chunk

    def get_microbatch_id(self, k: int) -> int:
        group_size = self.stages * self.num_model_chunks
        return (k // group_size) * self.stages + (k % self.stages)

    def _is_first(self, chunk: int) -> bool:
        return self.is_first_stage and chunk == 0

    def _is_last(self, chunk: int) -> bool:
        return self.is_last_stage and chunk == self.num_model_chunks - 1

    def _op_sequence(self) -> List[Tuple[str, int, int]]:
        ops: List[Tuple[str, int, int]] = []
        w, r = self.num_warmup_microbatches, self.num_microbatches_remaining
        for k in range(w):
            ops.append(("F", self.get_microbatch_id(k), self.get_model_chunk_id(k, True)))
        for k in range(r):
            fk = k + w
            ops.append(("F", self.get_microbatch_id(fk), self.get_model_chunk_id(fk, True)))
            ops.append(("B", self.get_microbatch_id(k), self.get_model_chunk_id(k, False)))
        for k in range(r, self.total_num_microbatches):
            ops.append(("B", self.get_microbatch_id(k), self.get_model_chunk_id(k, False)))
        return ops

    def steps(self):
        gb = not self.fused_fwd_bwd
        for kind, mb, chunk in self._op_sequence():
            cmds: List[object] = []
            if kind == "F":
                if not self._is_first(chunk):
                    cmds.append(ForwardPreprocessTask(mb, chunk, graph_break=not self.fused_send_recv))
                cmds.append(ForwardStepTask(mb, chunk, graph_break=gb))
                if not self._is_last(chunk):
                    cmds.append(ForwardPostprocessTask(mb, chunk, graph_break=True))
            else:
                if not self._is_last(chunk):
                    cmds.append(BackwardPreprocessTask(mb, chunk, graph_break=not self.fused_send_recv))
                cmds.append(BackwardStepTask(mb, chunk, graph_break=gb))
                if not self._is_first(chunk):
                    cmds.append(BackwardPostprocessTask(mb, chunk, graph_break=True))
            yield cmds
        yield [ReduceGradsTask()]


class TrainSchedule(PipeSchedule):
    """Clock-driven formulation of non-interleaved 1F1B (the reference keeps its deprecated predecessor,
    ``scheduler.py:545-677``, as a test oracle): one list per global clock tick, empty when the stage idles.  The
    non-empty ticks contain the same compute tasks in the same order as :class:`Train1F1BSchedule`."""

    def steps(self):
        n, pp, s = self.num_microbatches, self.stages, self.stage_id
        has_prev, has_next = self._valid_stage(self.prev_stage), self._valid_stage(self.next_stage)
        fwd_time = {}
        bwd_time = {}
        # event-driven simulation of all stages with unit-time ops gives the tick of every op on this stage
        seqs = [Train1F1BSchedule(n, pp, i)._op_sequence() for i in range(pp)]
        idx = [0] * pp
        done_f = [set() for _ in range(pp)]
        done_b = [set() for _ in range(pp)]
        tick = 0
        total = sum(len(q) for q in seqs)
        executed = 0
        while executed < total:
            ready = []
            for st in range(pp):
                if idx[st] >= len(seqs[st]):
                    continue
                kind, mb = seqs[st][idx[st]]
                if kind == "F":
                    ok = st == 0 or mb in done_f[st - 1]
                else:
                    ok = (st == pp - 1 and mb in done_f[st]) or (st < pp - 1 and mb in done_b[st + 1])
                if ok:
                    ready.append((st, kind, mb))
            assert ready, "1F1B schedule deadlocked"
            for st, kind, mb in ready:
                if st == s:
                    (fwd_time if kind == "F" else bwd_time)[mb] = tick
                idx[st] += 1
                executed += 1
            for st, kind, mb in ready:
                (done_f if kind == "F" else done_b)[st].add(mb)
            tick += 1
        by_tick = {}
        for mb, t in fwd_time.items():
            by_tick.setdefault(t, []).append(("F", mb))
        for mb, t in bwd_time.items():
            by_tick.setdefault(t, []).append(("B", mb))
        for t in range(tick):
            cmds: List[object] = []
            for kind, mb in by_tick.get(t, []):
                if kind == "F":
                    cmds += [ForwardPreprocessTask(mb), ForwardStepTask(mb)]
                    if has_next:
                        cmds.append(ForwardPostprocessTask(mb))
                else:
                    if has_next:
                        cmds.append(BackwardPreprocessTask(mb))
                    cmds.append(BackwardStepTask(mb))
                    if has_prev:
                        cmds.append(BackwardPostprocessTask(mb))
            if t == tick - 1:
                cmds.append(ReduceGradsTask())
            yield cmds
        if tick == 0:
            yield [ReduceGradsTask()]
