"""horovod_distributed_example_b200 - a B200-native data-parallel training framework with
the capabilities of the Horovod/Keras MNIST example (weikaolun/horovod-distributed-example).

    import horovod_distributed_example_b200 as hvd      # or: import hvd
    hvd.init(); hvd.size(); hvd.rank(); hvd.local_rank()
    opt = hvd.DistributedOptimizer(hvd.keras.optimizers.Adadelta(1.0 * hvd.size()))
    hvd.callbacks.BroadcastGlobalVariablesCallback(0)

See DESIGN.md for the architecture and INTEGRATION.md for the C ABI.
"""
from . import callbacks, keras  # noqa: F401
from .basics import (cross_rank, cross_size, device, gloo_built, gloo_enabled, init,  # noqa: F401
                     is_initialized, local_rank, local_size, mpi_built, mpi_enabled,
                     mpi_threads_supported, nccl_built, rank, shutdown, size)
from .collectives import (Adasum, Average, Compression, Sum, allgather, allgather_object,  # noqa: F401
                          allreduce, allreduce_, barrier, broadcast, broadcast_,
                          broadcast_object, broadcast_parameters, broadcast_variables,
                          grouped_allreduce_)
from .config import model_paths  # noqa: F401
from .optimizer import DistributedOptimizer  # noqa: F401

__version__ = "0.1.0"


def broadcast_global_variables(root_rank: int, model=None):
    """Function form of ``BroadcastGlobalVariablesCallback`` for a built model."""
    cb = callbacks.BroadcastGlobalVariablesCallback(root_rank)
    cb.set_model(model)
    cb.on_train_begin()
