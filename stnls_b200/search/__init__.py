"""stnls_b200.search -- mirrors lib/stnls/search/__init__.py for the hot-path searches."""
from . import non_local_search, refinement  # noqa: F401
from .api import extract_config, init  # noqa: F401
from .non_local_search import NonLocalSearch, NonLocalSearchFunction  # noqa: F401
from .refinement import RefineSearch, RefineSearchFunction  # noqa: F401
