"""B200-native batched rigid-body dynamics behind the GRiDCodeGenerator API."""
from .GRiDCodeGenerator import GRiDCodeGenerator  # noqa: F401
from .robot import URDFParser, available_robots, load_robot  # noqa: F401

__all__ = ["GRiDCodeGenerator", "URDFParser", "available_robots", "load_robot"]
