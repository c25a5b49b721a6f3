"""Process implementations of the reducer path, exported under the reference's names so that a
registry built by reflection (``inspect.getmembers(module, inspect.isfunction)``, reference
``core.py:97-103`` / ``tests/conftest.py:101-126``) finds the same functions."""

from .arrays import *  # noqa: F401,F403
from .comparison import *  # noqa: F401,F403
from .cubes import *  # noqa: F401,F403
from .logic import *  # noqa: F401,F403
from .math import *  # noqa: F401,F403
from .core import process, ParameterReference  # noqa: F401
from . import exceptions  # noqa: F401
