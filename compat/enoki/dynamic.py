"""enoki.dynamic stand-in (see enoki/__init__.py)."""
from . import make_types

Float32, Vector3f = make_types("cpu", False)
