from .model import (
    activation_function_selection,
    loss_function_selection,
    unsorted_segment_mean,
    unsorted_segment_sum,
    update_multibranch_heads,
    save_model,
    load_existing_model,
    load_existing_model_config,
    get_summary_writer,
    EarlyStopping,
    Checkpoint,
)
from .operations import get_edge_vectors_and_lengths
