from .train_validate_test import (
    get_head_indices,
    train_step,
    train,
    validate,
    test,
    train_validate_test,
)
