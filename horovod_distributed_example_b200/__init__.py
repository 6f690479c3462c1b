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


def load_model(filepath, custom_optimizers=None, custom_objects=None, compression=Compression.none):
    """``horovod.keras.load_model``: load a full-model ``.h5`` and wrap its optimizer in ``DistributedOptimizer``
    (state restored into the wrapped optimizer), ready to continue training.  Compiling with a distributed
    optimizer allocates the symmetric arena, which is a collective: EVERY rank must call this (as with
    Horovod), typically followed by ``BroadcastGlobalVariablesCallback(0)``."""
    from .utils import keras_h5
    return keras_h5.load_model(filepath, compile=True,
                               wrap_optimizer=lambda o: DistributedOptimizer(o, compression=compression))


def broadcast_optimizer_state(optimizer, root_rank: int = 0, model=None) -> None:
    """Make every rank's optimizer state equal to ``root_rank``'s.  Accepts this package's Keras-style optimizers
    (slot slabs live in the model's arena: pass ``model`` or rely on the optimizer's bound arena) and
    ``torch.optim`` optimizers (every tensor of ``state_dict()['state']`` plus the param-group scalars)."""
    import torch
    if size() == 1:
        return
    if isinstance(optimizer, torch.optim.Optimizer):
        sd = optimizer.state_dict()
        for _, st in sorted(sd["state"].items()):
            for k in sorted(st):
                if isinstance(st[k], torch.Tensor):
                    broadcast_(st[k], root_rank, name=f"opt.{k}")
        groups = broadcast_object([{k: v for k, v in g.items() if k != "params"} for g in sd["param_groups"]], root_rank)
        for g, new in zip(optimizer.param_groups, groups):
            g.update(new)
        return
    arena = model.arena if model is not None else getattr(optimizer, "_arena", None)
    if arena is None:
        raise ValueError("optimizer is not bound to a model yet: pass model=...")
    with torch.no_grad():
        for i, s in enumerate(arena.slots):
            broadcast_(s, root_rank, name=f"broadcast.slot{i}")
    it, lr = broadcast_object((optimizer.iterations, optimizer.lr), root_rank)
    optimizer.iterations = it
    optimizer.lr = lr


def start_timeline(file_path: str, mark_cycles: bool = False) -> None:
    """Start writing the Chrome-trace timeline at run time (same writer as ``HVD_TIMELINE``); rank 0 only."""
    from .basics import state
    from .utils.timeline import Timeline
    st = state()
    if st.rank == 0 and st.timeline is None:
        st.timeline = Timeline(file_path, mark_cycles=mark_cycles)


def stop_timeline() -> None:
    from .basics import state
    st = state()
    if st.timeline is not None:
        for hook in list(st.shutdown_hooks):        # engines drain their device-side records first
            hook()
        st.timeline.close()
        st.timeline = None


def ddl_built() -> bool:
    return False


def mlsl_built() -> bool:
    return False


def ccl_built() -> bool:
    return False


def cuda_built() -> bool:
    return True


def rocm_built() -> bool:
    return False


def broadcast_global_variables(root_rank: int, model=None):
    """Function form of ``BroadcastGlobalVariablesCallback`` for a built model."""
    cb = callbacks.BroadcastGlobalVariablesCallback(root_rank)
    cb.set_model(model)
    cb.on_train_begin()
