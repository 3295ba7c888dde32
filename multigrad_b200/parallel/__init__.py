"""Process bootstrap, communicators, collectives and the NVLink peer
workspace."""
from .comm import (Comm, init_distributed, get_world, is_distributed,  # noqa
                   local_device, shutdown)
from .collectives import (reduce_sum, split_subcomms,  # noqa: F401
                          split_subcomms_by_node, scatter_nd,
                          group_assignment)
