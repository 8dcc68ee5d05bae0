from .parallel import (DataParallel, EagerReducer, Group, ParallelEnv, assign_group_by_size, get_group, get_rank,  # noqa: F401
                       get_world_size, init_parallel_env, sync_params_buffers)
from .store import FileStore, TCPStore  # noqa: F401
from .fleet_utils import build_groups, fused_allreduce_gradients, fused_allreduce_gradients_with_group  # noqa: F401,E402
from .sharding import DygraphShardingOptimizer, GroupShardedOptimizerStage2, ShardingGroup  # noqa: F401,E402
