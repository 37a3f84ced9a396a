from .load_data import (
    dataset_loading_and_splitting,
    create_dataloaders,
    split_dataset,
    build_graphs,
    update_predicted_values,
    transform_raw_data_to_serialized,
)
