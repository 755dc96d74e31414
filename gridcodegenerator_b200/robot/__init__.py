"""Robot-model layer (URDFParser-compatible `robot` objects) and shipped robots."""
import os

from .model import Robot, Joint, Link, FixedJoint, LinkDesc, JointDesc
from .urdf import parse_urdf, parse_urdf_string, URDFParser
from . import floating  # noqa: F401  (pose / twist <-> six-joint-chain coordinates of a floating base)

LIBRARY_DIR = os.path.join(os.path.dirname(__file__), "library")
_CACHE = {}


def available_robots():
    return sorted(f[:-5] for f in os.listdir(LIBRARY_DIR) if f.endswith(".urdf"))


def load_robot(name: str) -> Robot:
    """Load a shipped robot description (iiwa14, hyq, atlas, pendulum2, chain3, branched6)."""
    if name not in _CACHE:
        path = os.path.join(LIBRARY_DIR, name + ".urdf")
        if not os.path.exists(path):
            raise KeyError(f"unknown robot {name!r}; available: {available_robots()}")
        _CACHE[name] = parse_urdf(path, name)
    return _CACHE[name]
