from .aggregate import *  # noqa: F401,F403
from .apply import *  # noqa: F401,F403
from .indices import *  # noqa: F401,F403
from .mask import *  # noqa: F401,F403
from .merge import *  # noqa: F401,F403
from .reduce import *  # noqa: F401,F403
from .cube import DataCube  # noqa: F401
