from .distributed import (
    GradientAverager,
    DataParallelModel,
    get_comm_size_and_rank,
    get_device,
    get_device_name,
    get_distributed_model,
    get_local_rank,
    is_model_distributed,
    nsplit,
    comm_reduce,
    setup_ddp,
)
