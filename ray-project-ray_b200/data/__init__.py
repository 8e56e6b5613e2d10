"""`ray.data`-shaped namespace: `from ray_project_ray_b200 import data as rd; rd.range(...)`."""
from .. import aggregate  # noqa: F401
from ..aggregate import AbsMax, AggregateFn, Count, Max, Mean, Min, Quantile, Std, Sum  # noqa: F401
from ..block import Block, BlockAccessor, BlockExecStats, BlockMetadata, DeviceBlock  # noqa: F401
from ..context import DataContext  # noqa: F401
from ..dataset import (Dataset, GroupedData, from_arrow, from_items, from_numpy, from_numpy_columns,  # noqa: F401
                       from_pandas, from_torch_columns, range)
