"""Pipeline schedules: exact task lists for small cases + a multi-stage replay that proves send/recv pairing,
ordering and deadlock freedom for pp in {2,4,8,16} x micro-batches in {1,4,8,32} (the sweep of the reference's
``test/unit_test/pipeline/test_scheduler.py:21-47``)."""
from __future__ import annotations

import itertools

import pytest

from neuronx_distributed_llama3_2_b200.pipeline.scheduler import (
    BackwardPostprocessTask, BackwardPreprocessTask, BackwardStepTask, ForwardPostprocessTask, ForwardPreprocessTask,
    ForwardStepTask, InferenceSchedule, ReduceGradsTask, Train1F1BSchedule, TrainInterleavedSchedule, TrainSchedule)


def test_1f1b_literal_last_stage():
    tasks = list(Train1F1BSchedule(num_microbatches=2, stages=4, stage_id=3).steps())
    assert tasks == [
        [ForwardPreprocessTask(0), ForwardStepTask(0)],
        [BackwardStepTask(0), BackwardPostprocessTask(0)],
        [ForwardPreprocessTask(1), ForwardStepTask(1)],
        [BackwardStepTask(1), BackwardPostprocessTask(1)],
        [ReduceGradsTask()],
    ]


def test_1f1b_literal_middle_stage_recv_bwd_before_send_fwd():
    tasks = list(Train1F1BSchedule(num_microbatches=3, stages=4, stage_id=2).steps())
    assert tasks == [
        [ForwardPreprocessTask(0), ForwardStepTask(0), ForwardPostprocessTask(0)],       # warm-up: send at once
        [ForwardPreprocessTask(1), ForwardStepTask(1)],                                  # steady: send deferred
        [BackwardPreprocessTask(0), ForwardPostprocessTask(1), BackwardStepTask(0), BackwardPostprocessTask(0)],
        [ForwardPreprocessTask(2), ForwardStepTask(2)],
        [BackwardPreprocessTask(1), ForwardPostprocessTask(2), BackwardStepTask(1), BackwardPostprocessTask(1)],
        [BackwardPreprocessTask(2), BackwardStepTask(2), BackwardPostprocessTask(2)],
        [ReduceGradsTask()],
    ]


def test_1f1b_first_stage_counts():
    s = Train1F1BSchedule(num_microbatches=8, stages=4, stage_id=0)
    assert (s.num_warmup_steps, s.num_steady_state_microbatches, s.num_remaining_microbatches) == (3, 5, 3)
    flat = [t for step in s.steps() for t in step]
    assert not any(isinstance(t, BackwardPostprocessTask) for t in flat)  # nothing to send backward from stage 0
    assert sum(isinstance(t, ForwardPostprocessTask) for t in flat) == 8
    assert isinstance(flat[-1], ReduceGradsTask)


def _replay(schedules, chunks=1, buffered=False):
    """Run all stages against each other and check FIFO pairing per channel, data dependencies and deadlock freedom.
    ``buffered=False``: BLOCKING rendezvous (a send completes only when the peer posts the matching receive) -- what
    the reference's 2-rank-collective send/recv needs and what 1F1B guarantees.  ``buffered=True``: sends are
    asynchronous (NCCL isend on a comm stream), receives block -- the contract of the interleaved schedule."""
    pp = len(schedules)
    chan = {}
    progs = [[t for step in s.steps() for t in step] for s in schedules]
    pc = [0] * pp
    fwd_done = [set() for _ in range(pp)]   # (mb, chunk) computed
    bwd_done = [set() for _ in range(pp)]
    recv_f = [set() for _ in range(pp)]
    recv_b = [set() for _ in range(pp)]

    def peer_of(stage, task):
        if isinstance(task, (ForwardPostprocessTask, BackwardPreprocessTask)):
            return (stage + 1) % pp
        return (stage - 1) % pp

    def dst_chunk(stage, task):
        c = task.model_chunk
        if isinstance(task, ForwardPostprocessTask):
            return c + 1 if stage == pp - 1 else c
        if isinstance(task, BackwardPostprocessTask):
            return c - 1 if stage == 0 else c
        return c

    progress = True
    while progress:
        progress = False
        for st in range(pp):
            while pc[st] < len(progs[st]):
                t = progs[st][pc[st]]
                key = (t.mb, t.model_chunk)
                if isinstance(t, ForwardStepTask):
                    first = st == 0 and t.model_chunk == 0
                    assert first or key in recv_f[st], f"stage {st} computes fwd {key} before receiving it"
                    fwd_done[st].add(key)
                elif isinstance(t, BackwardStepTask):
                    last = st == pp - 1 and t.model_chunk == chunks - 1
                    assert key in fwd_done[st]
                    assert last or key in recv_b[st], f"stage {st} computes bwd {key} before receiving the grad"
                    bwd_done[st].add(key)
                elif isinstance(t, ReduceGradsTask):
                    pass
                elif isinstance(t, (ForwardPostprocessTask, BackwardPostprocessTask)):
                    if isinstance(t, ForwardPostprocessTask):
                        assert key in fwd_done[st]
                    else:
                        assert key in bwd_done[st]
                    peer = peer_of(st, t)
                    want_type = ForwardPreprocessTask if isinstance(t, ForwardPostprocessTask) else BackwardPreprocessTask
                    if buffered:
                        chan.setdefault((st, peer, want_type), []).append((t.mb, dst_chunk(st, t)))
                        pc[st] += 1
                        progress = True
                        continue
                    if pc[peer] >= len(progs[peer]):
                        raise AssertionError(f"stage {st} sends {t} but stage {peer} has finished")
                    pt = progs[peer][pc[peer]]
                    if not (isinstance(pt, want_type) and (pt.mb, pt.model_chunk) == (t.mb, dst_chunk(st, t))):
                        break  # peer not at the matching receive yet: block
                    (recv_f if want_type is ForwardPreprocessTask else recv_b)[peer].add((pt.mb, pt.model_chunk))
                    pc[peer] += 1
                else:  # a receive: completes only through the sender's branch above
                    first_f = isinstance(t, ForwardPreprocessTask) and st == 0 and t.model_chunk == 0
                    last_b = isinstance(t, BackwardPreprocessTask) and st == pp - 1 and t.model_chunk == chunks - 1
                    if not (first_f or last_b):
                        if not buffered:
                            break
                        src = (st - 1) % pp if isinstance(t, ForwardPreprocessTask) else (st + 1) % pp
                        q = chan.get((src, st, type(t)), [])
                        if not q:
                            break
                        assert q[0] == key, f"stage {st} expects {key} but channel head is {q[0]}"
                        q.pop(0)
                        (recv_f if isinstance(t, ForwardPreprocessTask) else recv_b)[st].add(key)
                pc[st] += 1
                progress = True
    stuck = {st: progs[st][pc[st]] for st in range(pp) if pc[st] < len(progs[st])}
    assert not stuck, f"deadlock: {stuck}"
    return fwd_done, bwd_done


@pytest.mark.parametrize("pp,mb", list(itertools.product([2, 4, 8, 16], [1, 4, 8, 32])))
def test_1f1b_replay(pp, mb):
    fwd, bwd = _replay([Train1F1BSchedule(mb, pp, s) for s in range(pp)])
    for st in range(pp):
        assert fwd[st] == {(i, 0) for i in range(mb)} and bwd[st] == fwd[st]


@pytest.mark.parametrize("pp,mb", list(itertools.product([2, 4, 8], [1, 4, 8, 32])))
def test_clock_schedule_matches_1f1b_compute_order(pp, mb):
    for st in range(pp):
        def compute(sched):
            return [(type(t).__name__, t.mb) for step in sched.steps() for t in step
                    if isinstance(t, (ForwardStepTask, BackwardStepTask))]
        assert compute(TrainSchedule(mb, pp, st)) == compute(Train1F1BSchedule(mb, pp, st))
    # the clock formulation has one entry per global tick: all stages agree on the length
    assert len({len(list(TrainSchedule(mb, pp, st).steps())) for st in range(pp)}) == 1


@pytest.mark.parametrize("pp,chunks,mb", [(2, 2, 2), (2, 2, 4), (4, 2, 4), (4, 2, 8), (4, 3, 8), (8, 2, 16)])
def test_interleaved_replay(pp, chunks, mb):
    fwd, bwd = _replay([TrainInterleavedSchedule(mb, chunks, pp, s) for s in range(pp)], chunks=chunks, buffered=True)
    want = {(i, c) for i in range(mb) for c in range(chunks)}
    for st in range(pp):
        assert fwd[st] == want and bwd[st] == want


def test_interleaved_needs_multiple_of_pp():
    with pytest.raises(ValueError):
        TrainInterleavedSchedule(3, 2, 4, 0)


def test_inference_schedule():
    assert list(InferenceSchedule(2, 4, 1).steps()) == [
        [ForwardPreprocessTask(0), ForwardStepTask(0), ForwardPostprocessTask(0)],
        [ForwardPreprocessTask(1), ForwardStepTask(1), ForwardPostprocessTask(1)],
    ]
