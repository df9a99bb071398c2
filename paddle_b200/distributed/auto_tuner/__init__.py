from .tuner import AutoTuner  # noqa: F401
from .search import GridSearch  # noqa: F401
from .prune import prune_by_divisibility, prune_by_memory  # noqa: F401
from .recorder import HistoryRecorder  # noqa: F401
from .utils import default_candidates, search_all  # noqa: F401
