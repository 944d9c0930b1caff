"""Mirror of the reference package ``src.conv_onet`` (src/conv_onet/__init__.py)."""
from . import config, models

__all__ = ["config", "models"]
