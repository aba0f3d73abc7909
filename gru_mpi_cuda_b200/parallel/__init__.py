from .partition import Slice, partition, rank_info  # noqa: F401
