from .partition import Slice, partition, rank_info  # noqa: F401


def __getattr__(name):  # lazy: dp/symm import torch.distributed
    if name == "DPContext":
        from .dp import DPContext
        return DPContext
    if name == "TPContext":
        from .tp import TPContext
        return TPContext
    if name == "SymmBuffer":
        from .symm import SymmBuffer
        return SymmBuffer
    if name == "McBuffer":
        from .symm import McBuffer
        return McBuffer
    raise AttributeError(name)
