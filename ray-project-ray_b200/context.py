"""DataContext: the configuration singleton of the path.

Mirrors the flags of the reference's DataContext that the all-to-all path reads
(python/ray/data/context.py:138-330): `use_push_based_shuffle` (:51-53, :228),
`pipeline_push_based_shuffle_reduce_tasks` (:229), `target_shuffle_max_block_size` (:32, :223),
`use_polars` (:63, :235), free-form kv configs (:305-330), plus the switch that selects the
fused device scheduler.
"""
from __future__ import annotations

import os
import threading
from dataclasses import dataclass, field
from typing import Any, Dict, Optional

DEFAULT_SHUFFLE_TARGET_MAX_BLOCK_SIZE = 1024 * 1024 * 1024
MAX_SAFE_BLOCK_SIZE_FACTOR = 1.5

_lock = threading.Lock()
_current: Optional["DataContext"] = None


@dataclass
class DataContext:
    target_shuffle_max_block_size: int = DEFAULT_SHUFFLE_TARGET_MAX_BLOCK_SIZE
    use_push_based_shuffle: bool = bool(int(os.environ.get("RAY_DATA_PUSH_BASED_SHUFFLE", "0")))
    pipeline_push_based_shuffle_reduce_tasks: bool = True
    use_polars: bool = False  # accepted for API parity; the device path has one sort engine
    #: True: all-to-all ops run through GpuExchangeScheduler (one fused device pass per op).
    #: False: run the reference's task structure (map tasks, merge rounds, reduce tasks) over
    #: device blocks with the pull- or push-based scheduler, as `use_push_based_shuffle` says.
    use_fused_device_exchange: bool = bool(int(os.environ.get("B200X_FUSED_EXCHANGE", "1")))
    #: number of samples per block used to pick range-partition splitters
    #: (the reference takes 10 per block: sort_task_spec.py:159; sampling is cheap on device)
    sort_samples_per_block: int = 10
    warn_on_driver_memory_usage_bytes: Optional[int] = None
    _kv_configs: Dict[str, Any] = field(default_factory=dict)

    @staticmethod
    def get_current() -> "DataContext":
        global _current
        with _lock:
            if _current is None:
                _current = DataContext()
            return _current

    @staticmethod
    def _set_current(ctx: "DataContext") -> None:
        global _current
        with _lock:
            _current = ctx

    def get_config(self, key: str, default: Any = None) -> Any:
        return self._kv_configs.get(key, default)

    def set_config(self, key: str, value: Any) -> None:
        self._kv_configs[key] = value

    def remove_config(self, key: str) -> None:
        self._kv_configs.pop(key, None)
