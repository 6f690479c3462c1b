"""Short alias: ``import hvd`` == ``import horovod_distributed_example_b200``."""
import sys as _sys

import horovod_distributed_example_b200 as _pkg

_sys.modules[__name__] = _pkg
