"""Hooks run after (delayed) pipeline partitioning (reference
``src/neuronx_distributed/trainer/post_partition_hooks.py:5-35``).  Delayed tracing is hard-off in the reference
(``utils/model_utils.py:339-348``) and here, so the hooks execute immediately when registered after partition."""
from __future__ import annotations

from typing import Any, Callable, List, Tuple


class PostPartitionHooks:
    def __init__(self):
        self.hooks: List[Tuple[Callable, tuple, dict]] = []

    def register_post_partition_hook(self, callable_function: Callable, func_args=(), func_kwargs=None) -> None:
        if not callable(callable_function):
            raise TypeError("hook must be callable")
        self.hooks.append((callable_function, tuple(func_args), dict(func_kwargs or {})))

    def execute_all_hooks(self, model: Any = None) -> List[Any]:
        """Run the hooks in registration order, forget them, return their results."""
        pending, self.hooks = self.hooks, []
        return [fn(*args, **kwargs) for fn, args, kwargs in pending]


hooks = PostPartitionHooks()
