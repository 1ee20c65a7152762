"""Importable alias of the package directory ``rnnt-transducer_b200/`` (a hyphen cannot
appear in a Python module name). All code lives there; this file only extends the
package search path and re-exports the public API."""
import os as _os

__path__.insert(0, _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                                 "rnnt-transducer_b200"))

from ._api import *  # noqa: F401,F403,E402
from ._model import *  # noqa: F401,F403,E402
from ._data import *  # noqa: F401,F403,E402
from ._train import *  # noqa: F401,F403,E402
from ._api import __all__ as _api_all  # noqa: E402
from ._model import __all__ as _model_all  # noqa: E402
from ._data import __all__ as _data_all  # noqa: E402
from ._train import __all__ as _train_all  # noqa: E402

__all__ = list(_api_all) + list(_model_all) + list(_data_all) + list(_train_all)
