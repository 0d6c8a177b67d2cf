"""``neuronx_distributed.utils.timeline`` counterpart (reference ``utils/timeline.py:6-137``): the implementation lives
in ``pipeline/timeline.py`` (Chrome-trace events per pipeline task, gathered over the PP group at step end)."""
from ..pipeline.timeline import PPTimeline, Timeline  # noqa: F401


class Event:
    """One labelled interval on a rank's track (reference ``timeline.py:6-12``)."""

    def __init__(self, label: str, rank: int, start: int = -1, end: int = -1):
        self.label, self.rank, self.start, self.end = label, rank, start, end

    def as_chrome_event(self) -> dict:
        return {"name": self.label, "ph": "X", "pid": 0, "tid": self.rank, "ts": self.start,
                "dur": max(self.end - self.start, 0)}
