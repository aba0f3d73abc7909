from . import oracle  # noqa: F401
from .namegen import NameGenerator, decode  # noqa: F401
from .reference import TorchNameGen  # noqa: F401
