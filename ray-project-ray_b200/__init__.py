"""b200x: a B200-native implementation of Ray Data's all-to-all hot path
(Dataset.sort, random_shuffle, sort-based groupby().aggregate) behind the
reference's ExchangeTaskSpec / ExchangeTaskScheduler / BlockAccessor surface.

Import as `ray_project_ray_b200` (alias package at the repo root).
"""
__version__ = "0.1.0"
