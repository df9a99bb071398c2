from .group_sharded_optimizer_stage2 import GroupShardedOptimizerStage2  # noqa: F401
from .group_sharded_stage2 import GroupShardedStage2  # noqa: F401
from .group_sharded_stage3 import GroupShardedStage3  # noqa: F401
from .group_sharded_storage import GradStorage, InternalStorage, ParamStorage  # noqa: F401
from .group_sharded_utils import GroupShardedScaler  # noqa: F401
